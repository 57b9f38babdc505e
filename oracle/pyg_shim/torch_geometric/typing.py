"""TEST INFRASTRUCTURE ONLY."""
from typing import Optional, Tuple, Union
from torch import Tensor


class SparseTensor:  # never instantiated by the reference entry points
    pass


Adj = Union[Tensor, SparseTensor]
OptTensor = Optional[Tensor]
OptPairTensor = Tuple[Tensor, Optional[Tensor]]
Size = Optional[Tuple[int, int]]
