from .fc_stack import MLPStack, TrainStep, TransformerFCBlock, TransformerFCBlockSP  # noqa: F401
