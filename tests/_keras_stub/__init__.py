"""Test-only: a torch stand-in for `keras` (keras_torch.py) and a loader that runs the reference's own hot-path sources
unmodified on top of it (ref_loader.py).  Never imported by the product package or by bench.py."""
