"""Layer API mirror of kgcnn.layers for the hot path (same names, kwargs, call signatures)."""
