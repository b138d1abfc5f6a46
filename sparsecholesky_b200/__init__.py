"""spchol-b200: B200-native multifrontal sparse Cholesky behind the reference's host API."""
from .api import (  # noqa: F401
    CscMatrix, SChol, Expected, Sym,
    triplet_to_csc_matrix,
    triplet_to_csc_matrix_gpu, load_matrix_market_to_csc, load_matrix_market_to_csc_gpu, build_csc_matrix_from_pattern,
    etree, post_order, col_count, ereach, compute_levels, compute_supernodes, atree, order_amd, permute,
    schol, chol, chol_batched, residual, solve, lbuf_to_csc, csc_to_dense, device_count,
)
