"""Small, fast pass over every kernel family of libsfb200 for `compute-sanitizer` ("race
detection"):

    compute-sanitizer --tool memcheck  python tools/sanitize_smoke.py all
    compute-sanitizer --tool racecheck python tools/sanitize_smoke.py simt      # shared-memory hazards of the SIMT kernels
    compute-sanitizer --tool synccheck python tools/sanitize_smoke.py all

`all`  = conv fprop / stats / dgrad / fused dgrad+BatchNorm-reduction / wgrad on the smallest kernel cases, the stems,
         the BatchNorm / pooling / head / scores / Adam / staging kernels, and one tiny training step of the model;
`simt` = only the kernels that synchronise through __syncthreads / shuffles on ordinary shared memory (BatchNorm
         reductions, pooling, scores, Adam, staging): racecheck does not model mbarrier / TMA / tcgen05 ordering, so the
         tensor-core kernels are checked by memcheck + synccheck and by their parity tests instead.
Every result is still compared with its PyTorch reference (a sanitizer-clean wrong answer is not a pass).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

import kernel_cases as kc  # noqa: E402
import test_kernels_gpu as tk  # noqa: E402
from video_classification_b200 import _ffi  # noqa: E402


def simt_family():
    tk.test_bn_finalize_and_fold()
    tk.test_bn_add_relu(64, False)
    tk.test_bn_add_relu(256, True)
    tk.test_bn_bwd(64, True, False)
    tk.test_bn_bwd(256, True, True)
    tk.test_maxpool(64, 16, 16)
    tk.test_head_pool(20, 2, 2, (4, 2, 2))
    tk.test_fc()
    tk.test_video_scores_match_run_eval_tail(249, [7, 5, 10, 1, 3])
    tk.test_video_scores_ties_and_errors()
    tk.test_sparse_model_matches_reference_forward_and_gradients(33, 5, 249)
    tk.test_stage_clips_u8_against_torchvision_golden_and_fp32_layout()
    tk.test_layout_input_strided()
    tk.test_pack_weight_roundtrip()
    import test_adam_gpu as ta
    ta.test_adam_step_bit_exact_against_oracle(1e-2, 1.0)
    print("simt family OK", flush=True)


def tensor_family():
    small = [c for c in kc.CONV_CASES if c[0] in ("1x1x1_64_256", "1x3x3_64_64", "1x3x3_128_128_s2", "fast_1x3x3_16_16",
                                                   "lateral0_3x1x1_8_16_s4")]
    for case in small:
        for v in ("plain", "stats", "epi"):
            r = kc.run_conv_case(case, _ffi.IMPL_TC, v)
            assert r["err"] < 2e-2, (case[0], v, r)
        r = kc.run_dgrad_case(case, _ffi.IMPL_TC, accumulate=False)
        assert r["err"] < 2e-2, (case[0], "dgrad", r)
        r = kc.run_wgrad_case(case, _ffi.IMPL_TC)
        assert r["err"] < 3e-2, (case[0], "wgrad", r)
    for case in tk.DGRAD_BN_CASES[:3]:
        tk.test_conv_dgrad_bn_fused_reduction(case)
    print("tensor family OK", flush=True)


def model_step():
    import parity_utils as pu
    from oracle import slowfast_oracle as so
    _, product = pu.build_pair("REF", True)
    xs, xf = so.synthetic_clips(1, 8, 8, 64)
    xs, xf = xs.cuda(), xf.cuda()
    y = torch.tensor([3], device="cuda")
    product.train()
    for _ in range(2):
        product.zero_grad(set_to_none=True)
        loss = F.cross_entropy(product([xs, xf]), y)
        loss.backward()
    assert torch.isfinite(loss)
    product.eval()
    with torch.no_grad():
        assert torch.isfinite(product([xs, xf])).all()
    print("model step OK", flush=True)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    os.environ.setdefault("SFB_GRAPHS", "0")  # every launch goes through the stream, one by one
    torch.cuda.set_device(0)
    simt_family()
    if what == "all":
        tensor_family()
        model_step()
    torch.cuda.synchronize()
    print("sanitize_smoke done:", what, flush=True)
