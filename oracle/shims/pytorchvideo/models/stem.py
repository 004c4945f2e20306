from oracle.pv_restate import ResNetBasicStem, create_res_basic_stem  # noqa: F401
