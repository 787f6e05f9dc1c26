"""ORACLE support — test infrastructure, NOT the product.

Golden fixture for the RAdam update: runs the UNMODIFIED reference optimizer
(`/root/reference/radam.py`, class RAdam, built as train_fastspeech2.py:412-414 builds it) for 8 steps on a
seeded 1003-element tensor and stores the parameter after every step.  With beta2 = 0.98 steps 1-5 take the
degenerated-to-SGD branch (N_sma < 5) and steps 6-8 the rectified branch.  Build container only:

    python oracle/make_golden_radam.py
"""
from __future__ import annotations

import sys
import warnings
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from oracle import ref_loader  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"
N, STEPS = 1003, 8


def inputs():
    gen = torch.Generator().manual_seed(77)
    p0 = torch.randn(N, generator=gen)
    grads = torch.randn(STEPS, N, generator=gen) * torch.logspace(-3, 0, N)  # gradient scales over 3 decades
    return p0, grads


def main():
    if ref_loader.REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    import radam  # the reference's own file, imported in place

    p0, grads = inputs()
    p = torch.nn.Parameter(p0.clone())
    opt = radam.RAdam([p], lr=1e-3, betas=(0.9, 0.98), eps=1e-9)
    out = []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # deprecated add_/addcmul_ overloads in radam.py:58-59
        for t in range(STEPS):
            p.grad = grads[t].clone()
            opt.step()
            out.append(p.detach().clone())
    st = opt.state[p]
    np.savez_compressed(GOLDEN / "radam_steps.npz", p0=p0.numpy(), grads=grads.numpy(),
                        p_steps=torch.stack(out).numpy(), exp_avg=st["exp_avg"].numpy(),
                        exp_avg_sq=st["exp_avg_sq"].numpy())
    print("wrote", GOLDEN / "radam_steps.npz")


if __name__ == "__main__":
    main()
