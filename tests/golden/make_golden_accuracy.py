"""Golden vectors of utils.sequence_accuracy (utils.py:44-80) from the LIVE reference function.  utils.py imports
prettytable (absent in the build container), so that one module is stubbed; the function itself runs
unmodified.  Run in the build container only:  python tests/golden/make_golden_accuracy.py"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.modules.setdefault("prettytable", types.SimpleNamespace(PrettyTable=object))
sys.path.insert(0, "/root/reference")
import utils as ref_utils  # noqa: E402


def main():
    out = {}
    for i, (B, T, V, pad, seed) in enumerate([(4, 6, 50, 3, 0), (7, 9, 333, 0, 1), (3, 5, 17, 3, 2)]):
        g = torch.Generator().manual_seed(seed)
        preds = torch.randn(B, T, V, generator=g)
        targets = torch.randint(0, V, (B, T), generator=g)
        targets[:, -2:] = pad                                  # trailing padding like generate_json_data.py:76
        # make a share of the targets likely hits so that both k give non-trivial values
        hit = torch.rand(B, T, generator=g) < 0.4
        boost = torch.zeros(B, T, V).scatter_(2, targets.unsqueeze(-1), 3.0) * hit.unsqueeze(-1)
        preds = preds + boost
        out[f"case{i}_cfg"] = np.array([B, T, V, pad, seed], dtype=np.int64)
        out[f"case{i}_preds"] = preds.numpy()
        out[f"case{i}_targets"] = targets.numpy()
        out[f"case{i}_acc"] = np.array([ref_utils.sequence_accuracy(preds, targets, k, ignore_index=pad) for k in (1, 3, 5)],
                                       dtype=np.float64)
    # no non-padding token at all -> 0 (utils.py:80)
    out["empty_acc"] = np.array([ref_utils.sequence_accuracy(torch.randn(2, 3, 9), torch.zeros(2, 3, dtype=torch.long), 1)],
                                dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "acc_sequence_accuracy.npz"), **out)
    print({k: v.tolist() for k, v in out.items() if k.endswith("_acc")})


if __name__ == "__main__":
    main()
