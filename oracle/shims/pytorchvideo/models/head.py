from oracle.pv_restate import ResNetBasicHead, create_res_basic_head, create_res_roi_pooling_head  # noqa: F401
