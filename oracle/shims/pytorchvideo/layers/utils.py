from oracle.pv_restate import set_attributes  # noqa: F401
