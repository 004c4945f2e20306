from oracle.pv_restate import (BottleneckBlock, ResBlock, ResStage, create_bottleneck_block,  # noqa: F401
                               create_res_block, create_res_stage)
