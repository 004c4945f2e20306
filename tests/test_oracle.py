"""CPU tests of the oracle (no GPU): the self-contained restatement (oracle/slowfast_oracle.py, the one that travels
to the GPU box) against (a) the committed golden fixtures, which were produced by the reference's own
model/my_slowfast.py imported verbatim, and (b) -- where /root/reference exists -- that verbatim import itself."""
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import ref_import, slowfast_oracle as so

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
torch.set_num_threads(max(1, min(8, os.cpu_count() or 1)))


def _build(meta):
    torch.manual_seed(meta["seed"])
    m = so.build_oracle(num_class=249, fuse=True, **so.VARIANTS[meta["variant"]])
    so.randomize_bn_(m, meta["bn_seed"], meta["final_gamma"])
    so.disable_dropout_(m)
    return m


@pytest.mark.parametrize("name", ["ref_small", "hyb_small"])
def test_restated_oracle_reproduces_golden(name):
    fx = torch.load(os.path.join(GOLDEN, name + ".pt"))
    meta = fx["meta"]
    m = _build(meta)
    assert list(m.state_dict().keys()) == fx["state_dict_keys"]
    assert len(fx["state_dict_keys"]) == 734 and meta["n_params"] == 38077321
    chk = sum(float(v.double().abs().sum()) for v in m.state_dict().values() if v.dtype.is_floating_point)
    assert abs(chk - meta["weight_checksum"]) <= 1e-9 * meta["weight_checksum"]
    xs, xf = so.synthetic_clips(meta["N"], meta["Ts"], meta["Tf"], meta["S"], seed=meta["clip_seed"])
    m.eval()
    with torch.no_grad():
        ev = m([xs, xf])
    assert torch.allclose(ev, fx["eval_logits"], rtol=1e-4, atol=1e-4 * float(fx["eval_logits"].abs().max()))
    if name == "hyb_small":
        return  # train-mode HYB at 224^2 is covered on the GPU; keep the CPU suite short
    m.train()
    out = m([xs, xf])
    loss = F.cross_entropy(out, torch.tensor(meta["labels"]))
    loss.backward()
    assert torch.allclose(out, fx["train_logits"], rtol=1e-4, atol=1e-5)
    assert abs(float(loss) - fx["loss"]) < 1e-4
    got = {n: p.grad for n, p in m.named_parameters() if p.grad is not None}
    assert set(got) == set(fx["grad_norm"])
    dead = [n for n, p in m.named_parameters() if p.grad is None]
    assert len(dead) == 48 and all(".residual." in n or ".res_unit." in n for n in dead)
    for n, g in got.items():
        ref_n = fx["grad_norm"][n]
        assert abs(float(g.norm()) - ref_n) <= 2e-3 * ref_n + 1e-9, n
        s = g.flatten()[fx["grad_idx"][n]]
        assert float((s - fx["grad_samples"][n]).norm()) <= 5e-3 * float(fx["grad_samples"][n].norm()) + 1e-9, n


@pytest.mark.skipif(not ref_import.available(), reason="/root/reference only exists in the authoring container")
@pytest.mark.parametrize("variant", ["REF", "HYB"])
def test_restated_oracle_is_bit_identical_to_verbatim_reference(variant):
    torch.manual_seed(0)
    ref = ref_import.build_reference_model(249, True, (5, 15), (64, 8), variant=variant)
    torch.manual_seed(0)
    mine = so.build_oracle(**so.VARIANTS[variant])
    sd_r, sd_m = ref.state_dict(), mine.state_dict()
    assert list(sd_r.keys()) == list(sd_m.keys())
    for k in sd_r:
        assert torch.equal(sd_r[k], sd_m[k]), k  # same init, same RNG consumption order
    for m in (ref, mine):
        so.randomize_bn_(m, 5, 0.1)
        so.disable_dropout_(m)
    S, Ts, Tf = (64, 8, 8) if variant == "REF" else (224, 8, 32)
    xs, xf = so.synthetic_clips(1 if variant == "HYB" else 2, Ts, Tf, S)
    ref.eval()
    mine.eval()
    with torch.no_grad():
        assert torch.equal(ref([xs, xf]), mine([xs, xf]))
    if variant == "REF":
        ref.train()
        mine.train()
        y = torch.tensor([3, 77])
        F.cross_entropy(ref([xs, xf]), y).backward()
        F.cross_entropy(mine([xs, xf]), y).backward()
        for (n, a), (_, b) in zip(ref.named_parameters(), mine.named_parameters()):
            assert (a.grad is None) == (b.grad is None), n
            if a.grad is not None:
                assert torch.equal(a.grad, b.grad), n


@pytest.mark.skipif(not ref_import.available(), reason="/root/reference only exists in the authoring container")
def test_reference_rejects_8_32_frames_like_survey_says():
    """the reference graph as written cannot take slow 8 / fast 32 frames (torch.cat size mismatch)."""
    torch.manual_seed(0)
    ref = ref_import.build_reference_model(249, True, (5, 15), (64, 8), variant="REF").eval()
    with pytest.raises(RuntimeError, match="Sizes of tensors must match"):
        with torch.no_grad():
            ref([torch.randn(1, 5, 8, 64, 64), torch.randn(1, 15, 32, 64, 64)])


def test_bf16_twin_keeps_the_graph_and_only_moves_storage_points():
    torch.manual_seed(0)
    m = so.build_oracle()
    so.randomize_bn_(m, 1, 0.1)
    so.disable_dropout_(m)
    so.round_conv_weights_to_bf16_(m)
    xs, xf = so.synthetic_clips(1, 4, 4, 64)
    m.eval()
    with torch.no_grad():
        a = m([xs, xf])
        handles = so.add_bf16_storage_hooks_(m)
        b = m([xs, xf])
        for h in handles:
            h.remove()
        c = m([xs, xf])
    assert torch.equal(a, c)
    err = float((a - b).abs().max() / a.abs().max())
    assert 0 < err < 2e-2


def test_no_fuse_variant_and_fuse_flag():
    torch.manual_seed(0)
    m = so.build_oracle(fuse=False)
    assert not any("multipathway_fusion" in k for k in m.state_dict())
    xs, xf = so.synthetic_clips(1, 4, 4, 64)
    m.eval()
    with torch.no_grad():
        assert m([xs, xf]).shape == (1, 249)


# ---- input staging oracle against torchvision's own transforms ------------------------------------
def _staging_golden():
    return torch.load(os.path.join(os.path.dirname(__file__), "golden", "staging_small.pt"))


def test_staging_oracle_reproduces_torchvision_golden():
    from oracle import staging_oracle as st
    g = _staging_golden()
    assert torch.equal(st.stage_clips(g["frames"]), g["plain"])
    assert torch.equal(st.stage_clips(g["frames"], g["windows"], g["padding"]), g["aug"])
    ramp = torch.arange(256, dtype=torch.uint8).reshape(1, 256, 1)
    assert torch.equal(st.to_tensor_normalize(ramp)[0, 0], g["lut"])


def test_staging_oracle_split_and_s2d_layout():
    from oracle import staging_oracle as st
    g = _staging_golden()
    x = st.stage_clips(g["frames"])
    xs, xf = st.prepare_slowfast_data(x, slow_t_stride=2)
    assert xs.shape == (3, 5, 1, 20, 20) and xf.shape == (3, 15, 2, 20, 20)
    assert torch.equal(xs[:, :, 0], x[:, 0, 0:5]) and torch.equal(xf[:, :, 1], x[:, 1, 5:20])
    s2d = st.space_to_depth_bf16(xf.contiguous(), 64)
    assert s2d.shape == (3, 2, 10, 10, 64) and float(s2d[..., 60:].abs().max()) == 0.0
    # channel (sh*2+sw)*C + c of s2d pixel (h, w) is input channel c of pixel (2h+sh, 2w+sw)
    assert torch.equal(s2d[1, 1, 3, 4, 2 * 15 + 7], xf[1, 7, 1, 7, 8].to(torch.bfloat16))
    assert torch.equal(s2d[2, 0, 9, 0, 1 * 15 + 14], xf[2, 14, 0, 18, 1].to(torch.bfloat16))


def test_raw_clips_host_checks():
    from video_classification_b200.staging import RawClips, random_crop_offsets
    off = random_crop_offsets(500, 64, torch.Generator().manual_seed(0))
    assert off.dtype == torch.int32 and off.shape == (500, 2) and int(off.min()) == -6 and int(off.max()) == 6
    rc = RawClips(torch.zeros((2, 8, 4, 4, 21), dtype=torch.uint8), slow_t_stride=4)
    rc.validate()
    assert rc.pathways() == ((0, 5, 4), (5, 15, 1))
    with pytest.raises(ValueError):
        RawClips(torch.zeros((2, 6, 4, 4, 21), dtype=torch.uint8), slow_t_stride=4).validate()
    with pytest.raises(TypeError):
        RawClips(torch.zeros((2, 8, 4, 4, 21))).validate()


# ---- evaluation-tail oracle ------------------------------------------------------------------
def test_eval_oracle_hand_example_and_sparse_model_keys():
    import numpy as np
    from oracle import eval_oracle as eo
    from video_classification_b200 import evaluation as ev
    logits = np.log(np.array([[0.7, 0.2, 0.1], [0.1, 0.8, 0.1], [0.6, 0.3, 0.1], [0.2, 0.2, 0.6]], dtype=np.float32))
    r = eo.run_eval_tail(logits, np.array([0, 0, 0, 2]), [3, 1])
    assert np.allclose(r["ps"].sum(1), 1.0, atol=1e-6)
    assert np.allclose(r["video_mean"][0], [1.4 / 3, 1.3 / 3, 0.1], atol=1e-6)
    assert r["video_top1"].tolist() == [0, 2] and r["acc"] == 1.0
    # product SparseModel: same constructor, parameters and state_dict keys as the reference module; CUDA only
    ref, mine = eo.SparseModelRef(7, 3), ev.SparseModel(7, 3)
    assert list(ref.state_dict().keys()) == list(mine.state_dict().keys()) == \
        [f"fcs.{c}.{n}" for c in range(7) for n in ("weight", "bias")]
    mine.load_state_dict(ref.state_dict(), strict=True)
    from video_classification_b200 import _ffi
    with pytest.raises(_ffi.SfbError):
        mine(torch.rand(2, 3, 7))
