import torch.nn as nn


class BaseModel(nn.Module):
    """audiotools.ml.BaseModel: the decode path only reads `.device`."""

    @property
    def device(self):
        return next(self.parameters()).device
