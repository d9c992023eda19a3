from .kernel import Kernel
from .matern import Matern
from .squared_exponential import SquaredExponential
from .kernel_params import GPParams

__all__ = ["Kernel", "Matern", "SquaredExponential", "GPParams"]
