def generate_structured_noise_batch_vectorized(*a, **k):  # SDE-BrushNet only (magicdrive_stdit3.py:18)
    raise NotImplementedError("oracle shim")
