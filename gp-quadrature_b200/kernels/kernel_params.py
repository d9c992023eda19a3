"""Log-parameterised hyperparameter container (mirror of the reference's kernels/kernel_params.py:9-55).

`raw` holds log(lengthscale, variance, ..., noise variance) as one nn.Parameter in the current
torch default dtype; the kernel reads its hypers back through this object.
"""
import torch
from torch import nn


class GPParams(nn.Module):
    def __init__(self, kernel, init_sig2):
        super().__init__()
        self.kernel = kernel
        if hasattr(kernel, "iter_hypers"):
            pairs = list(kernel.iter_hypers())
            self.hypers_names = [n for n, _ in pairs]
            vals = [float(v) for _, v in pairs]
        else:
            self.hypers_names = []
            vals = []
        vals.append(float(init_sig2))
        self.raw = nn.Parameter(torch.log(torch.tensor(vals, dtype=torch.get_default_dtype())))
        if hasattr(kernel, "_gp_params_ref"):
            kernel._gp_params_ref = self

    @property
    def pos(self):
        return self.raw.exp()

    @property
    def sig2(self):
        return self.pos[-1]
