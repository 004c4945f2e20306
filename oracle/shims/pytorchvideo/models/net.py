from oracle.pv_restate import DetectionBBoxNetwork, MultiPathWayWithFuse, Net  # noqa: F401
