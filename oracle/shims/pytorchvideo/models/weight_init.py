from oracle.pv_restate import init_net_weights  # noqa: F401
