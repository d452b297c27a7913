from .fc_stack import (MLPStack, TrainStep, TransformerFCBlock, TransformerFCBlockMegatron,  # noqa: F401
                       TransformerFCBlockSP, TransformerFCStack)
