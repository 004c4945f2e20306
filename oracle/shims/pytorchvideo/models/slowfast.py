from oracle.pv_restate import (FastToSlowFusionBuilder, FuseFastToSlow, PoolConcatPathway,  # noqa: F401
                               create_slowfast)
