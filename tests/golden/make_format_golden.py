"""Generate tests/golden/format_*.npz by IMPORTING the reference's own packer
(/root/reference/scripts/model_convert/gptq2marlin.py: get_perms, marlin_repack_qweight,
marlin_permute_scales).  Run in the build container only (the reference tree does not exist on
the GPU box); the .npz files are committed.

    python tests/golden/make_format_golden.py
"""
import importlib.util
import os
import sys

import numpy as np
import torch

REF = "/root/reference/scripts/model_convert/gptq2marlin.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_ref():
    spec = importlib.util.spec_from_file_location("ref_gptq2marlin", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = load_ref()
    perm, scale_perm, scale_perm_single = ref.get_perms()
    np.savez_compressed(os.path.join(HERE, "format_perms.npz"),
                        perm=perm.numpy().astype(np.int32),
                        scale_perm=np.array(scale_perm, dtype=np.int32),
                        scale_perm_single=np.array(scale_perm_single, dtype=np.int32))
    rng = np.random.default_rng(20261104)
    cases = {}
    # (K, N): smallest legal block, one group, multi-chunk, ragged-ish N (=64*3), a 4096-ish slice
    for k, n in [(16, 64), (32, 128), (128, 64), (256, 192), (512, 256), (1024, 320)]:
        gptq = rng.integers(-2**31, 2**31, size=(k // 8, n), dtype=np.int64).astype(np.int32)
        marlin = ref.marlin_repack_qweight(torch.from_numpy(gptq), 4, k, n).numpy()
        cases[f"gptq_{k}x{n}"] = gptq
        cases[f"marlin_{k}x{n}"] = marlin
        if k % 128 == 0:
            s = rng.uniform(0.002, 0.02, size=(k // 128, n)).astype(np.float16)
            sp = ref.marlin_permute_scales(torch.from_numpy(s), k, n, 128).numpy()
            cases[f"scales_{k}x{n}"] = s
            cases[f"scales_marlin_{k}x{n}"] = sp
        sc = rng.uniform(0.002, 0.02, size=(1, n)).astype(np.float16)
        scp = ref.marlin_permute_scales(torch.from_numpy(sc), k, n, -1).numpy()
        cases[f"chan_scales_{k}x{n}"] = sc
        cases[f"chan_scales_marlin_{k}x{n}"] = scp
    np.savez_compressed(os.path.join(HERE, "format_cases.npz"), **cases)
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    sys.exit(main())
