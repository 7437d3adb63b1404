"""Model families the reference's examples train: the NYC-taxi MLP,
DLRM, and the large MLP of the trainer benchmark."""
from raydp_b200.models.big_mlp import BigMLP  # noqa: F401
from raydp_b200.models.dlrm import DLRM, DLRMSpec  # noqa: F401
from raydp_b200.models.mlp import NYCTaxiMLP, TabularMLP  # noqa: F401
