"""BaseModel (reference src/models/base_model.py:9-470): style injected by concatenation
(`Decoder` / `DecoderConcat`).  Shares the AdaINModel schedule; the BaseModel-specific decoders
are the next row of the scope table and are not built yet."""
from .model import Model


class BaseModel(Model):
    def __init__(self, args):
        super().__init__(args)
        raise NotImplementedError(
            "BaseModel's Decoder/DecoderConcat/StyleEncoder are not on the accelerated path yet "
            "(SURVEY.md §8f-2); use --model AdaINModel")

    def set_inputs(self, inputs):
        raise NotImplementedError

    def optimize_parameters(self, global_iter):
        raise NotImplementedError
