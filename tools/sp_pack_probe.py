"""Round 2: the SP collide on the packed FP32x2 pipe of sm_100 (FADD2 / FMUL2 / FFMA2, two sites per issue slot) against the
scalar build, per vector width and register budget.  Library variants are separate builds selected by LBM_B200_LIBDIR
(inside the package so that they travel to the box):
    make -C latticeboltzmann_b200/csrc OUT=../exp/nopack EXTRA=-DLBM_NO_F32X2 ../exp/nopack/liblbm_b200_sp.so
    make -C latticeboltzmann_b200/csrc OUT=../exp/minb6 EXTRA=-DLBM_AA_MINB=6 ../exp/minb6/liblbm_b200_sp.so      (80 registers)
    make -C latticeboltzmann_b200/csrc OUT=../exp/minb5 EXTRA=-DLBM_AA_MINB=5 ../exp/minb5/liblbm_b200_sp.so      (96 registers)
usage: sp_pack_probe.py            (parent: runs itself once per variant)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = os.path.join(ROOT, "latticeboltzmann_b200")


def child(tag):
    from latticeboltzmann_b200 import Lattice
    for nx, ny, T in ((400, 4000, 2000), (800, 4000, 1000), (8192, 32768, 64)):
        with Lattice(nx, ny, "sp", inplace=True) as lat:
            lat.InitializeArrays()
            for vec in (4, 2):
                for block in (128, 256):
                    lat.set_tuning(vec=vec, block=block)
                    lat.run_timed(T // 4)
                    ms = min(lat.run_timed(T) for _ in range(2))
                    print(json.dumps(dict(variant=tag, nx=nx, ny=ny, vec=vec, block=block, mlups=round(nx * ny * T / ms / 1e3, 1),
                                          checksum=lat.checksum())), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        child(sys.argv[1])
    else:
        for tag, d in (("packed_64reg", PKG), ("scalar_64reg", os.path.join(PKG, "exp", "nopack")),
                       ("packed_80reg", os.path.join(PKG, "exp", "minb6")), ("packed_96reg", os.path.join(PKG, "exp", "minb5"))):
            if not os.path.exists(os.path.join(d, "liblbm_b200_sp.so")):
                print(json.dumps(dict(variant=tag, missing=d)), flush=True)
                continue
            subprocess.run([sys.executable, os.path.abspath(__file__), tag], env={**os.environ, "LBM_B200_LIBDIR": d}, check=False)
