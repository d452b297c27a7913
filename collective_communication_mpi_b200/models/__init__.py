from .fc_stack import MLPStack, TrainStep, TransformerFCBlock  # noqa: F401
