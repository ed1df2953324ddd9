"""Makes `3dod_b200/` a sys.path root: `tdod` (library binding + batched API) and the drop-in `ProposalNetwork`
mirror are imported as top-level packages, exactly as the reference's callers import `ProposalNetwork.*`."""
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
