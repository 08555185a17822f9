"""Where the two big in-loop GEMMs (gate GEMM of the forward step, dz GEMM of the backward step) spend their time inside the
REAL train step: CTA (0,0,0) leaves %globaltimer stamps of its phases (sat_gemm_phase_stamps); this prints the deltas of the
last launch of each, averaged over a few steps.  Usage (GPU box): python tools/gemm_phases.py > gpurun_out/gemm_phases.txt"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402

PHASES = ["entry", "after PDL wait", "first stage landed", "last MMA issued", "accumulator complete", "split-K exchange done",
          "epilogue done"]


def main():
    from sat_b200 import _lib, synth
    from sat_b200.decoder import Decoder
    from sat_b200.trainer import FusedTrainStep
    CFG = bench.CFG
    dev = torch.device("cuda", 0)
    L = _lib.lib()
    L.sat_gemm_phase_stamps.argtypes = [C.c_int, C.POINTER(C.c_uint64)]
    dec = Decoder(CFG["V"], CFG["D"], tf=True, attention=True, compute_dtype=torch.bfloat16)
    dec.load_state_dict(synth.make_params(CFG["V"], CFG["D"], CFG["H"], seed=42))
    dec = dec.to(dev).train()
    tr = FusedTrainStep(dec, lr=1e-4, alpha_c=1.0)
    host = bench.make_host_batches(2, CFG["B"])
    batches = [(a.to(dev), c.to(dev)) for a, c in host]
    for i in range(4):
        tr.step(*batches[i % 2])
    torch.cuda.synchronize()
    _lib.check(L.sat_gemm_phase_stamps(1, None))
    acc = [[0.0] * 8 for _ in range(2)]
    n = 5
    for i in range(n):
        tr.step(*batches[i % 2])
        out = (C.c_uint64 * 16)()
        _lib.check(L.sat_gemm_phase_stamps(1, out))
        for g in range(2):
            for j in range(7):
                acc[g][j] += (out[g * 8 + j] - out[g * 8]) / 1e3 / n
    _lib.check(L.sat_gemm_phase_stamps(0, None))
    for g, name in enumerate(("gate GEMM (EpiLstm, 32x2 tiles, split-K 2)", "dz GEMM (EpiDzGate, 40x2 tiles)")):
        print(f"# {name}: us since kernel entry of CTA (0,0,0), mean of {n} steps (last launch of each step)")
        prev = 0.0
        for j in range(7):
            if j == 5 and acc[g][j] <= 0:
                continue
            print(f"  {PHASES[j]:26s} {acc[g][j]:8.2f}   (+{acc[g][j] - prev:6.2f})")
            prev = acc[g][j]


if __name__ == "__main__":
    main()
