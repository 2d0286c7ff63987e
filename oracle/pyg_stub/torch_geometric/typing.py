from typing import Optional, Tuple, Union

from torch import Tensor


class SparseTensor:  # placeholder: only named in type hints by the reference
    pass


Adj = Union[Tensor, SparseTensor]
OptTensor = Optional[Tensor]
PairTensor = Tuple[Tensor, Tensor]
OptPairTensor = Tuple[Tensor, Optional[Tensor]]
