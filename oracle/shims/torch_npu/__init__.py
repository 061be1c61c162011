"""Empty stand-in: blocks.py:23-24 imports torch_npu when CUDA is absent."""
