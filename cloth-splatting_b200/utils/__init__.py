"""Mirror of the reference's `utils` package, restricted to what the hot path needs (loss_utils)."""
