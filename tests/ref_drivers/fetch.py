"""Copies the reference's DRIVER files (the callers of the hot path, which this repo does not rebuild) from
/root/reference into tests/ref_drivers/_ref/ so that tests/test_drivers_unchanged.py can run them, unmodified, on
the GPU box, where /root/reference does not exist.  The copies are test fixtures like oracle/_ref/: git-ignored
(never part of the history), not gpurun-ignored (they travel with the snapshot).  Run by __graft_entry__.build()
whenever /root/reference is present; nothing in the product imports them.

Usage: python tests/ref_drivers/fetch.py [reference_root]"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
FILES = [
    "train_model.py", "test_model.py", "image_dataloaders.py", "parser_config.py", "visualization.py",
    "models/__init__.py", "models/fetch_network.py", "models/baseline_net.py", "models/efficient_net.py",
    "models/efficient_net_v2.py", "models/perceiver.py",
    "configs/example_config.py",
]


def fetch(ref_root="/root/reference"):
    if not os.path.isdir(ref_root):
        return False
    for rel in FILES:
        src = os.path.join(ref_root, rel)
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
    return True


if __name__ == "__main__":
    ok = fetch(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    print("fetched" if ok else "reference not present", DEST)
