"""B200-native `local-store` (LocalAI's built-in vector store), sm_100a only.

`Store` mirrors the reference's Go `Store`
(/root/reference/backend/go/stores/store.go) on top of the C ABI in
include/tfp_store.h.  The CUDA library is the only implementation: importing
the package works anywhere, but using a store needs the built library and a
B200 -- there is no CPU fallback.
"""
from .store import Store, StoreError, FindContext  # noqa: F401
from ._lib import lib_path, load  # noqa: F401
