"""Subprocess-isolated kernel diagnostics for the B200 box: prints the error of every conv case for both
implementations, one process per kind so a device trap in one kind does not hide the others.
Usage: python tools/kernel_diag.py [kind ...]   (kinds: fprop fused dgrad wgrad)"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def child(kind):
    import torch
    import kernel_cases as kc
    from video_classification_b200 import _ffi
    for impl_name, impl in (("simt", _ffi.IMPL_SIMT), ("tc", _ffi.IMPL_TC)) if kind == "wgrad" else (("tc", _ffi.IMPL_TC),):
        for case in kc.CONV_CASES:
            variants = {"fprop": ["plain"], "fused": ["pro", "epi", "stats"], "dgrad": ["d", "dacc"], "wgrad": ["w", "wpro"]}[kind]
            for v in variants:
                if kind == "dgrad" and case[0].startswith("stem"):
                    continue
                t0 = time.time()
                try:
                    if kind in ("fprop", "fused"):
                        r = kc.run_conv_case(case, impl, v)
                    elif kind == "dgrad":
                        r = kc.run_dgrad_case(case, impl, accumulate=(v == "dacc"))
                    else:
                        ci_real = {"stem_slow_8_64": 5, "stem_fast_16_8": 15, "stem_slow_16_64": 5}.get(case[0])
                        r = kc.run_wgrad_case(case, impl, pro=(v == "wpro"), ci_real=ci_real)
                    flag = "OK " if r["err"] < 2e-2 else "BAD"
                    print(f"{flag} {kind:6s} {impl_name:4s} {v:5s} {case[0]:28s} " +
                          " ".join(f"{k}={val:.3e}" for k, val in r.items()) + f"  ({time.time()-t0:.2f}s)", flush=True)
                except Exception as e:  # a CUDA trap is sticky: report and stop this process
                    print(f"EXC {kind:6s} {impl_name:4s} {v:5s} {case[0]:28s} {type(e).__name__}: {str(e)[:300]}", flush=True)
                    if "CUDA" in str(e) or "cuda" in str(e):
                        return 1
    return 0


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        sys.exit(child(sys.argv[2]))
    kinds = sys.argv[1:] or ["fprop", "fused", "dgrad", "wgrad"]
    for k in kinds:
        print(f"==== {k} ====", flush=True)
        try:
            subprocess.run([sys.executable, os.path.abspath(__file__), "--child", k], timeout=600)
        except subprocess.TimeoutExpired:
            print(f"TIMEOUT {k}", flush=True)
