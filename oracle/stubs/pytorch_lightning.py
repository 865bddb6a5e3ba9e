"""TEST INFRASTRUCTURE ONLY. Minimal stand-in for `pytorch_lightning` so that the reference's
`lightning/xunet.py` and `lightning/diff3d.py` can be imported on a box without Lightning.
Only what those two files touch: `LightningModule` (an nn.Module with a `.device` property,
reference use: lightning/xunet.py:309,322,327) and a dummy `Trainer`."""
import torch


class LightningModule(torch.nn.Module):
    @property
    def device(self):
        for p in self.parameters():
            return p.device
        return torch.device("cpu")

    def optimizers(self):  # lightning/diff3d.py:82,124 -- only reached from training_step/warmup
        return self._oracle_optimizer


class Trainer:  # never used by the oracle
    def __init__(self, *a, **k):
        raise RuntimeError("oracle stub: pytorch_lightning.Trainer is not available")
