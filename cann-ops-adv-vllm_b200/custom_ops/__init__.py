"""Drop-in replacement of the reference's `custom_ops` package (extension/custom_ops/__init__.py):
same operator names and positional signatures, bound to hand-written sm_100a CUDA kernels."""
from .add_custom import (cent_select, compute_cent, incre_flash_attention_v4, select_position,
                         sparse_paged_attention, sparse_paged_fusion_attention)
from .add_custom import FusedDecodePlan  # noqa: F401  (extension: launch-ready fused executor)
from .add_custom import assign_page_clusters  # noqa: F401  (extension: centroid maintenance)

__all__ = ["incre_flash_attention_v4", "sparse_paged_attention", "compute_cent", "select_position",
           "cent_select", "sparse_paged_fusion_attention", "FusedDecodePlan", "assign_page_clusters"]
