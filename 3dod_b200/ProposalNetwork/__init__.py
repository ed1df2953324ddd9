"""Drop-in mirror of the reference's `ProposalNetwork` package for its proposal-and-scoring hot path
(SURVEY.md 8(b)): same module paths, function names, argument meaning and error behaviour as
/root/reference/ProposalNetwork/{proposals,scoring,utils}, every function forwarding to libtdod.so (hand-written
CUDA for sm_100a) through `tdod.api`.  Put `3dod_b200/` ahead of the reference on sys.path and the callers in
cubercnn/modeling/roi_heads/roi_heads.py:2-4,30-34 import these instead.

Inputs must be CUDA tensors: there is no CPU path (a CPU tensor raises ValueError).  Pieces of the reference
package that are not on the hot path (plotting helpers, iou_3d, the autograd hull of the weak loss,
convex_outline) are deliberately absent -- see DESIGN.md "out of scope"."""
