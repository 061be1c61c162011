"""TEST INFRASTRUCTURE — copies the handful of pure-Python files of the REAL reference that its denoiser and samplers
import (16 files, 270 KB; list below = what `oracle/ref_loader.py` pulls into sys.modules) from /root/reference into
the git-ignored `oracle/_ref/`, so that the UNMODIFIED reference travels to the GPU box with the snapshot, like the
built libmdv2.so does: there it is the CPU baseline of bench.py (`cpu_baseline.kind = "reference"`) and the
`--impl reference` arm.  Nothing under `oracle/_ref/` is committed, and the product never imports it.

Run in the build container:  python oracle/populate_ref.py      (also called by __graft_entry__.build())
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("MDV2_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = """
magicdrivedit/acceleration/__init__.py
magicdrivedit/acceleration/checkpoint.py
magicdrivedit/acceleration/communications.py
magicdrivedit/acceleration/parallel_states.py
magicdrivedit/models/layers/__init__.py
magicdrivedit/models/layers/blocks.py
magicdrivedit/models/magicdrive/__init__.py
magicdrivedit/models/magicdrive/embedder.py
magicdrivedit/models/magicdrive/magicdrive_stdit3.py
magicdrivedit/models/magicdrive/utils.py
magicdrivedit/models/vae/utils.py
magicdrivedit/models/vae/vae_temporal.py
magicdrivedit/registry.py
magicdrivedit/schedulers/rf/__init__.py
magicdrivedit/schedulers/rf/rectified_flow.py
magicdrivedit/utils/misc.py
magicdrivedit/utils/inference_utils.py
""".split()


def populate(verbose=True) -> bool:
    if not os.path.isdir(os.path.join(SRC, "magicdrivedit")):
        if verbose:
            print(f"populate_ref: {SRC} not present (GPU box?) — keeping whatever oracle/_ref holds")
        return os.path.isdir(os.path.join(DST, "magicdrivedit"))
    for f in FILES:
        d = os.path.join(DST, f)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(os.path.join(SRC, f), d)
    if verbose:
        print(f"populate_ref: {len(FILES)} reference files -> {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if populate() else 1)
