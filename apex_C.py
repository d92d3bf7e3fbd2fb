"""Drop-in for apex's ``apex_C`` extension (flatten / unflatten).  The DDP
wrapper in this repository does not use them (gradients are gathered straight into persistent
buckets), but the names stay importable for callers of the old API."""
from torch._utils import _flatten_dense_tensors as flatten  # noqa: F401
from torch._utils import _unflatten_dense_tensors as unflatten  # noqa: F401
