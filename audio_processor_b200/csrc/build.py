"""Build the native library.

    python -m audio_processor_b200.csrc.build            # libafe.so  (nvcc, sm_100a) -- the product
    python -m audio_processor_b200.csrc.build --emu      # tests/_emu/libafe_emu.so (g++, TEST ONLY)

The emulation build compiles the very same kernel sources as plain C++ with AFE_EMULATE and
runs each block program phase by phase on the host; it exists so kernel index arithmetic
can be checked against the oracle without a GPU.  The package never loads it.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
SOURCES = ["afe_runtime.cu", "afe_host.cc", "afe_tables.cc", "afe_resample.cu", "afe_fbank.cu", "afe_acoustic.cu"]
HEADERS = ["afe_platform.h", "afe_fft.h", "afe_rfft512.h", "afe_tables.h", "afe_ctx.h", "afe_fbank_kernels.h",
           "afe_fbank_pipe.h", "afe_resample_kernels.h", "afe_acoustic_kernels.h", "afe_dct_gen.h", "afe_mel_slots.h", os.path.join(ROOT, "include", "afe.h")]
CUDA_LIB = os.path.join(PKG, "libafe.so")
EMU_LIB = os.path.join(ROOT, "tests", "_emu", "libafe_emu.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--threads", "0"]


def _stale(target: str) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    deps = [os.path.join(HERE, s) for s in SOURCES] + [h if os.path.isabs(h) else os.path.join(HERE, h) for h in HEADERS]
    deps.append(os.path.abspath(__file__))
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build_cuda(force: bool = False, verbose: bool = False, extra=(), out: str = CUDA_LIB) -> str:
    """``out`` / ``extra``: development variants (tools/kernel_lab.py builds them under tools/_bin/)."""
    if not force and not _stale(out):
        return out
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, *NVCC_FLAGS, *extra, "-shared", "-o", out] + [os.path.join(HERE, s) for s in SOURCES]
    if verbose:
        cmd.insert(1, "-Xptxas"); cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libafe.so")
    return out


def build_emu(force: bool = False) -> str:
    if not force and not _stale(EMU_LIB):
        return EMU_LIB
    os.makedirs(os.path.dirname(EMU_LIB), exist_ok=True)
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DAFE_EMULATE", "-ffp-contract=off", "-o", EMU_LIB] + os.environ.get("AFE_EMU_EXTRA", "").split() + os.environ.get("AFE_EMU_EXTRA", "").split()
    for s in SOURCES:
        cmd += ["-x", "c++", os.path.join(HERE, s)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("g++ failed building libafe_emu.so")
    return EMU_LIB


if __name__ == "__main__":
    if "--emu" in sys.argv:
        print(build_emu(force=True))
    else:
        print(build_cuda(force=True, verbose="-v" in sys.argv))
