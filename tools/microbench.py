#!/usr/bin/env python
"""Runs drtvam_b200/csrc/microbench.cu on cuda:0 and prints one JSON object:
fp32 reduction-atomic throughput into a 1 GiB (> L2) and a 64 MiB (L2-resident)
array for three address patterns, REDG.v4, and shared-memory accumulate rates.
These are the atomic-roofline denominators SURVEY.md section 8(d) asks for."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from drtvam_b200 import build  # noqa: E402


def main():
    build.build()
    lib = C.CDLL(build.MICRO)
    lib.drtvam_mb_red.restype = C.c_double
    lib.drtvam_mb_red.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_int, C.c_int, C.c_int]
    lib.drtvam_mb_smem.restype = C.c_double
    lib.drtvam_mb_smem.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
    torch.cuda.set_device(0)
    out = {}
    blocks = 148 * 16
    for label, log2n in (("1GiB", 28), ("64MiB", 24)):
        n = 1 << log2n
        buf = torch.zeros(n + 32 * 512 + 64, device="cuda", dtype=torch.float32)
        for pname, pattern, stride in (("warp_contiguous", 0, 0), ("warp_stride400", 1, 400), ("random", 2, 0)):
            g = lib.drtvam_mb_red(buf.data_ptr(), n - 1, pattern, stride, 256, blocks, 0)
            out[f"red_f32_{pname}_{label}_Gadds"] = round(g, 2)
        g = lib.drtvam_mb_red(buf.data_ptr(), n - 1, 0, 0, 256, blocks, 1)
        out[f"red_v4_f32_{label}_Gadds"] = round(g, 2)
        del buf
    o = torch.zeros(blocks, device="cuda")
    for name, mode in (("atoms_int", 0), ("atoms_f32_cas", 1), ("lds_fadd_sts", 2)):
        out[f"smem_{name}_Gops"] = round(lib.drtvam_mb_smem(o.data_ptr(), mode, 4096, 148 * 4), 2)
    lib.drtvam_mb_smem2.restype = C.c_double
    lib.drtvam_mb_smem2.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    for name, mode in (("red_f32", 0), ("red_u32", 1), ("cas64_f32x2", 2), ("lds_fadd_sts", 3), ("lds", 4)):
        for share in (1, 2, 4):
            for threads in (256, 1024):
                g = lib.drtvam_mb_smem2(o.data_ptr(), mode, 8192, 148 * (4 if threads == 256 else 2), threads, share)
                out[f"smem2_{name}_share{share}_t{threads}_Gops"] = round(g, 2)
    # legacy tensor path (mma.sync m16n8k8 tf32 -> HMMA): dense TFLOP/s by warps per SM and accumulator tiles per warp
    lib.drtvam_mb_mma_tf32.restype = C.c_double
    lib.drtvam_mb_mma_tf32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    for threads in (128, 256, 512, 1024):
        for ilp in (1, 4, 16):
            t = lib.drtvam_mb_mma_tf32(o.data_ptr(), ilp, 4096, 148, threads)
            out[f"mma_sync_tf32_m16n8k8_w{threads // 32}_ilp{ilp}_TFLOPs"] = round(t, 1)
    # the instruction of the split-fp16 patch gather: mma.sync m16n8k8 f16 with fp32 accumulators
    lib.drtvam_mb_mma_f16.restype = C.c_double
    lib.drtvam_mb_mma_f16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    for threads in (128, 256, 512):
        for ilp in (1, 4, 16):
            t = lib.drtvam_mb_mma_f16(o.data_ptr(), ilp, 4096, 148, threads)
            out[f"mma_sync_f16_m16n8k8_w{threads // 32}_ilp{ilp}_TFLOPs"] = round(t, 1)
    lib.drtvam_mb_mma_f16k16.restype = C.c_double
    lib.drtvam_mb_mma_f16k16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
    for threads in (128, 256, 512):
        for ilp in (1, 4, 16):
            t = lib.drtvam_mb_mma_f16k16(o.data_ptr(), ilp, 4096, 148, threads)
            out[f"mma_sync_f16_m16n8k16_w{threads // 32}_ilp{ilp}_TFLOPs"] = round(t, 1)
    torch.cuda.synchronize()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
