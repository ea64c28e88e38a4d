"""Builds the in-tree CUDA libraries (sm_100a) and the C driver.  nvcc cross-compiles without a GPU."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def build(verbose=False):
    out = None if verbose else subprocess.DEVNULL
    subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j2", "all"], check=True, stdout=out)
    subprocess.run(["make", "-C", os.path.join(ROOT, "driver"), "all"], check=True, stdout=out)


if __name__ == "__main__":
    build(verbose=True)
