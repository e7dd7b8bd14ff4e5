"""the rest of Synthesizer.infer (emb_g, TextEncoder, z_p sampling, reverse flow) in front of the decoder.
CPU tests pin the oracle restatement (oracle/synth_oracle.py) and the module's seeded initialisation against the golden
generated from the unmodified reference (oracle/make_golden_r2.py --synth); GPU tests run rvc_b200.Synthesizer.infer
(module -> ctypes -> rvcsyn_infer) against that golden and against the oracle at a larger, ragged shape."""
import os

import numpy as np
import pytest
import torch

from oracle import nsf_hifigan_oracle as O
from oracle import synth_oracle as SO
from helpers import GOLDEN, record_parity

SYN_48K = [1025, 32, 192, 192, 768, 2, 6, 3, 0, "1", [3, 7, 11], [[1, 3, 5], [1, 3, 5], [1, 3, 5]], [12, 10, 2, 2], 512,
           [24, 20, 4, 4], 109, 256, 48000]
CFGD = dict(inter_channels=192, hidden_channels=192, n_heads=2, n_layers=6, kernel_size=3)


def _golden():
    return np.load(os.path.join(GOLDEN, "synth_48k_B2_T40.npz"))


def _module(d):
    """rvc_b200.Synthesizer initialised under the fixture's seed, with the fixture's non-zero flow.post tensors."""
    import rvc_b200

    torch.manual_seed(int(d["init_seed"]))
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval()
    with torch.no_grad():
        for k in d.files:
            if k.startswith("sd::"):
                dict(m.named_parameters())[k[4:]].copy_(torch.from_numpy(d[k]))
    return m


def _inputs(d):
    return (torch.from_numpy(d["phone"].astype(np.float32)), torch.from_numpy(d["lengths"]), torch.from_numpy(d["pitch"]),
            torch.from_numpy(d["nsff0"]), torch.from_numpy(d["sid"]), torch.from_numpy(d["noise_zp"]),
            torch.from_numpy(d["noise_src"].astype(np.float32)))


def test_module_init_and_oracle_match_the_reference_golden():
    d = _golden()
    m = _module(d)
    sd = m.state_dict()
    cks = [k for k in d.files if k.startswith("ck::")]
    # every tensor infer.py:484-485 keeps (the fixture was written after `del net_g.enc_q`)
    assert sorted(k[4:] for k in cks) == sorted(k for k in sd.keys() if not k.startswith("enc_q."))
    for k in cks:
        v = sd[k[4:]].double()
        np.testing.assert_allclose(np.array([float(v.sum()), float(v.abs().sum())]), d[k], rtol=1e-12, atol=1e-12, err_msg=k)
    phone, lens, pitch, nsff0, sid, nzp, nsrc = _inputs(d)
    W = SO.folded({k: v for k, v in sd.items() if not k.startswith("enc_q.")})
    zm, g, mask, z_p, m_p, logs_p = SO.synthesizer_front(W, CFGD, phone, lens, pitch, sid, nzp)
    assert float((z_p - torch.from_numpy(d["z_p"])).abs().max()) <= 2e-5
    assert float((m_p - torch.from_numpy(d["m_p"])).abs().max()) <= 2e-5
    assert float((zm - torch.from_numpy(d["z"])).abs().max()) <= 2e-5
    assert float(zm[1, :, 29:].abs().max()) == 0.0  # item 1 is 29 frames long: masked frames are exactly zero
    Wd = O.folded_weights({k[4:]: v for k, v in sd.items() if k.startswith("dec.")})
    y = O.decoder_forward(Wd, O.CONFIG_48K, zm, nsff0, g, nsrc)
    assert float((y - torch.from_numpy(d["o"])).abs().max()) <= 2e-5
    with pytest.raises(RuntimeError):
        m.infer(phone, lens, pitch, nsff0, sid)  # CPU tensors: no fallback


@pytest.mark.skipif(__import__("oracle.stage_ref", fromlist=["ref_path"]).ref_path() is None, reason="reference not importable")
def test_seeded_init_equals_the_live_reference_including_enc_q():
    import sys
    from oracle.stage_ref import ref_path
    import rvc_b200

    if ref_path() not in sys.path:
        sys.path.insert(0, ref_path())
    from rvc.lib.algorithm.synthesizers import Synthesizer as Ref

    torch.manual_seed(11)
    a = Ref(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").state_dict()
    torch.manual_seed(11)
    b = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").state_dict()
    assert list(a.keys()) == list(b.keys()) and len(a) == 560
    assert all(torch.equal(a[k], b[k]) for k in a)
    torch.manual_seed(12)
    a = Ref(*SYN_48K, use_f0=0, text_enc_hidden_dim=768, vocoder="HiFi-GAN").state_dict()
    torch.manual_seed(12)
    b = rvc_b200.Synthesizer(*SYN_48K, use_f0=0, text_enc_hidden_dim=768, vocoder="HiFi-GAN").state_dict()
    assert list(a.keys()) == list(b.keys()) and all(torch.equal(a[k], b[k]) for k in a)


def test_flip_folding_is_a_channel_reversal():
    """The library folds the flow's Flip layers into the neighbouring 1x1 convolutions (rvcsyn.cu): a coupling layer that
    sees flip(x) reads the upper channel half in reverse order and updates the lower half in reverse order."""
    torch.manual_seed(0)
    x = torch.randn(1, 8, 5)
    pre, post = torch.randn(6, 4, 1), torch.randn(4, 6, 1)
    xf = torch.flip(x, [1])
    x0, x1 = xf[:, :4], xf[:, 4:]
    m = torch.nn.functional.conv1d(torch.tanh(torch.nn.functional.conv1d(x0, pre)), post)
    ref = torch.flip(torch.cat([x0, x1 - m], 1), [1])                       # back in the storage frame
    pre_f, post_f = torch.flip(pre, [1]), torch.flip(post, [0])             # reversed input / output channels
    m2 = torch.nn.functional.conv1d(torch.tanh(torch.nn.functional.conv1d(x[:, 4:], pre_f)), post_f)
    got = torch.cat([x[:, :4] - m2, x[:, 4:]], 1)
    assert float((got - ref).abs().max()) <= 1e-6


@pytest.mark.gpu
def test_synthesizer_infer_matches_the_reference_golden():
    d = _golden()
    dev = torch.device("cuda", 0)
    m = _module(d).to(dev)
    phone, lens, pitch, nsff0, sid, nzp, nsrc = _inputs(d)
    o, x_mask, (z, _, _, _) = m.infer(phone.to(dev), lens.to(dev), pitch.to(dev), nsff0.to(dev), sid.to(dev),
                                      noise_zp=nzp.to(dev), noise_src=nsrc.to(dev), mode="fp32")
    o_ref, z_ref = torch.from_numpy(d["o"]), torch.from_numpy(d["z"])
    assert o.shape == o_ref.shape == (2, 1, 40 * 480) and x_mask.shape == (2, 1, 40)
    ez, eo = float((z.cpu() - z_ref).abs().max()), float((o.cpu() - o_ref).abs().max())
    assert float(z[1, :, 29:].abs().max()) == 0.0
    assert ez <= 1e-4, ez
    assert eo <= 2e-3, eo
    of, _, _ = m.infer(phone.to(dev), lens.to(dev), pitch.to(dev), nsff0.to(dev), sid.to(dev),
                       noise_zp=nzp.to(dev), noise_src=nsrc.to(dev), mode="fast")
    corr = O.multires_logmel_corr(of.cpu(), o_ref, 48000)
    record_parity({"test": "synthesizer_infer_golden_B2_T40", "z_max_abs": ez, "fp32_max_abs": eo, "fast_logmel": corr,
                   "launches": m.last_launch_count})
    assert corr["min"] >= 0.999, corr


@pytest.mark.gpu
def test_synthesizer_infer_against_the_oracle_at_a_ragged_batch():
    """B = 3, T = 333 (several attention tiles, a partial last tile), lengths 333 / 200 / 64; default noise = device
    generator draws in the reference's order."""
    import rvc_b200

    dev = torch.device("cuda", 0)
    torch.manual_seed(77)
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval()
    with torch.no_grad():
        for i in range(4):
            m.flow.flows[2 * i].post.weight.normal_(0, 0.05)
            m.flow.flows[2 * i].post.bias.normal_(0, 0.05)
    sd = m.state_dict()
    B, T = 3, 333
    gen = torch.Generator().manual_seed(5)
    phone = torch.randn(B, T, 768, generator=gen)
    pitch = torch.randint(1, 256, (B, T), generator=gen)
    nsff0 = O.synth_f0(B, T, gen)
    sid = torch.tensor([0, 57, 108])
    lens = torch.tensor([333, 200, 64])
    m = m.to(dev)
    torch.manual_seed(9)
    o, x_mask, (z, _, _, _) = m.infer(phone.to(dev), lens.to(dev), pitch.to(dev), nsff0.to(dev), sid.to(dev), mode="fp32")
    torch.manual_seed(9)
    nzp = torch.randn(B, 192, T, device=dev).cpu()
    torch.rand(1, 1, 1, device=dev)
    nsrc = torch.randn(B, T * 480, 1, device=dev).cpu().reshape(B, -1)
    W = SO.folded({k: v for k, v in sd.items() if not k.startswith("enc_q.")})
    zm, g, mask, *_ = SO.synthesizer_front(W, CFGD, phone, lens, pitch, sid, nzp)
    Wd = O.folded_weights({k[4:]: v for k, v in sd.items() if k.startswith("dec.")})
    y = O.decoder_forward(Wd, O.CONFIG_48K, zm, nsff0, g, nsrc)
    ez, eo = float((z.cpu() - zm).abs().max()), float((o.cpu() - y).abs().max())
    record_parity({"test": "synthesizer_infer_oracle_B3_T333_ragged", "z_max_abs": ez, "fp32_max_abs": eo})
    assert torch.equal(x_mask.cpu(), mask)
    assert ez <= 2e-4, ez
    assert eo <= 2e-3, eo


@pytest.mark.gpu
def test_tensor_core_front_end_against_the_fp32_front_end():
    """mode="fast": q|k|v, o, FFN, proj and the coupling layers' convolutions on the tensor-core back-end from fp16 operand
    copies (fp32 accumulate, bias, residual, LayerNorm, attention, gate).  Ragged batch (lengths 333 / 200 / 64): z of
    the fast front end against the fp32 front end (option front_fp32) with the same injected noise - rows beyond an
    item's length exactly zero, everything else within the operand rounding - and the waveform at the fast bar."""
    import rvc_b200

    dev = torch.device("cuda", 0)
    torch.manual_seed(78)
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval()
    with torch.no_grad():
        for i in range(4):
            m.flow.flows[2 * i].post.weight.normal_(0, 0.05)
            m.flow.flows[2 * i].post.bias.normal_(0, 0.05)
    B, T = 3, 333
    gen = torch.Generator().manual_seed(6)
    phone = torch.randn(B, T, 768, generator=gen).to(dev)
    pitch = torch.randint(1, 256, (B, T), generator=gen).to(dev)
    nsff0 = O.synth_f0(B, T, gen).to(dev)
    sid = torch.tensor([0, 57, 108]).to(dev)
    lens = torch.tensor([333, 200, 64]).to(dev)
    nzp = torch.randn(B, 192, T, generator=gen).to(dev)
    nsrc = torch.randn(B, T * 480, generator=gen).to(dev)
    m = m.to(dev)
    kw = dict(noise_zp=nzp, noise_src=nsrc)
    o32, _, (z32, _, _, _) = m.infer(phone, lens, pitch, nsff0, sid, mode="fp32", **kw)
    n32 = m.last_launch_count
    h = m._ensure_handle(dev)
    h.set_option("front_fp32", 1)
    o_a, _, (z_a, _, _, _) = m.infer(phone, lens, pitch, nsff0, sid, mode="fast", **kw)
    h.set_option("front_fp32", 0)
    o_b, _, (z_b, _, _, _) = m.infer(phone, lens, pitch, nsff0, sid, mode="fast", **kw)
    o_b2, _, (z_b2, _, _, _) = m.infer(phone, lens, pitch, nsff0, sid, mode="fast", **kw)
    assert torch.equal(z_a, z32)                      # fp32 front end whatever the decoder's mode
    assert torch.equal(z_b, z_b2) and torch.equal(o_b, o_b2)
    assert float(z_b[1, :, 200:].abs().max()) == 0.0 and float(z_b[2, :, 64:].abs().max()) == 0.0
    ez = float((z_b - z32).abs().max())
    rel = float((z_b - z32).norm() / z32.norm())
    corr = O.multires_logmel_corr(o_b.cpu()[:1], o32.cpu()[:1], 48000)
    record_parity({"test": "tensor_core_front_end_B3_T333_ragged", "z_max_abs_vs_fp32_front": ez, "z_rel_l2": rel, "fast_logmel_vs_fp32": corr,
                   "launches_fast": m.last_launch_count, "launches_fp32": n32})
    assert rel <= 5e-3, (rel, ez)
    assert ez <= 0.1, ez
    assert corr["min"] >= 0.999, corr


def _ref_attention(qkv, ek, ev, lens, heads=2, W=10, round16=False):
    """fp64 restatement of attentions.py:89-125 on a (B, 3H, T) q | k | v tensor (optionally rounded to fp16 first, as the
    tensor-core kernel reads it)."""
    import math

    B, C3, T = qkv.shape
    H = C3 // 3
    kc = H // heads
    f = (lambda x: x.half().double()) if round16 else (lambda x: x.double())
    q, k, v = [f(qkv[:, i * H:(i + 1) * H].reshape(B, heads, kc, T).transpose(2, 3)) for i in range(3)]
    ek, ev = f(ek), f(ev)
    idx = torch.arange(T)
    d = idx[None, :] - idx[:, None]
    rel = q @ ek.t()  # (B, heads, T, 2W+1)
    s = q @ k.transpose(2, 3) + torch.where(d.abs() <= W, rel.gather(3, (d.clamp(-W, W) + W).expand(B, heads, T, T)),
                                            torch.zeros((), dtype=torch.float64))
    s = s / math.sqrt(kc)
    for b, n in enumerate(lens):
        dead = torch.ones(T, T, dtype=torch.bool)
        dead[:n, :n] = False
        s[b, :, dead] = -1e4
    p = torch.softmax(s, -1)
    pb = torch.zeros(B, heads, T, 2 * W + 1, dtype=torch.float64)
    for j in range(2 * W + 1):
        t = torch.arange(max(0, W - j), min(T, T - (j - W)))
        pb[:, :, t, j] = p[:, :, t, t + (j - W)]
    return (p @ v + pb @ ev).transpose(2, 3).reshape(B, H, T)


@pytest.mark.gpu
@pytest.mark.parametrize("B,T,lens", [(3, 333, [333, 200, 64]), (2, 128, [128, 1]), (1, 1000, [777])])
def test_tensor_core_attention_against_the_fp32_kernel(B, T, lens):
    """att_mma.cu (scores and P V as tcgen05 GEMMs over 128 x 128 tiles, online softmax, relative keys through a 21-column
    side GEMM, relative values as one more K = 32 MMA into the same accumulator) against the fp32 attention kernel on the
    SAME fp16-rounded q | k | v: output of layer 0's attention ("att" tap), ragged lengths, partial last tiles, a one-frame
    item.  The kernel is held to an fp64 restatement on the fp16-rounded q | k | v it reads (5e-4 of the output scale); the
    distance to the fp32 kernel is the rounding of q | k | v itself (seeded random weights make the logits large: 1-3 %),
    bounded here and, end to end, by z of the whole front end with either kernel."""
    import rvc_b200

    dev = torch.device("cuda", 0)
    torch.manual_seed(79)
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval()
    with torch.no_grad():  # relative embeddings large enough to matter
        for l in range(6):
            m.enc_p.encoder.attn_layers[l].emb_rel_k.normal_(0, 0.3)
            m.enc_p.encoder.attn_layers[l].emb_rel_v.normal_(0, 0.3)
    ek = m.enc_p.encoder.attn_layers[0].emb_rel_k.detach()[0].clone()
    ev = m.enc_p.encoder.attn_layers[0].emb_rel_v.detach()[0].clone()
    gen = torch.Generator().manual_seed(7)
    phone = torch.randn(B, T, 768, generator=gen).to(dev)
    pitch = torch.randint(1, 256, (B, T), generator=gen).to(dev)
    nsff0 = O.synth_f0(B, T, gen).to(dev)
    sid = torch.zeros(B, dtype=torch.long, device=dev)
    lens_t = torch.tensor(lens).to(dev)
    nzp = torch.randn(B, 192, T, generator=gen).to(dev)
    nsrc = torch.randn(B, T * 480, generator=gen).to(dev)
    m = m.to(dev)
    kw = dict(noise_zp=nzp, noise_src=nsrc, mode="fast")
    m.infer(phone, lens_t, pitch, nsff0, sid, **kw)
    h = m._ensure_handle(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    att, z = {}, {}

    def tap(name, C, stop):
        h.set_stop(stop)
        m.infer(phone, lens_t, pitch, nsff0, sid, **kw)
        dst = torch.empty(B, C, T, dtype=torch.float32, device=dev)
        assert h.debug_read(name, dst, m._workspace, B, T, stream) == dst.numel()
        torch.cuda.synchronize()
        h.set_stop(0)
        return dst.cpu()

    try:
        h.set_option("att_mma", 0)
        qkv = tap("qkv", 576, 2)
        for on in (0, 1):
            h.set_option("att_mma", on)
            att[on] = tap("att", 192, 3)
            _, _, (zz, _, _, _) = m.infer(phone, lens_t, pitch, nsff0, sid, **kw)
            z[on] = zz.cpu()
    finally:
        h.set_stop(0)
        h.set_option("att_mma", 1)
    assert bool(torch.isfinite(att[1]).all())
    r_exact, r_16 = _ref_attention(qkv, ek, ev, lens), _ref_attention(qkv, ek, ev, lens, round16=True)
    e_kernel, e_round, e_fp32 = [], [], []
    for b, n in enumerate(lens):  # rows beyond an item's length are zeroed downstream (LayerNorm mask): compare the live rows
        sc = float(r_exact[b, :, :n].abs().max()) + 1e-9
        e_kernel.append(float((att[1][b, :, :n].double() - r_16[b, :, :n]).abs().max()) / sc)
        e_round.append(float((att[1][b, :, :n] - att[0][b, :, :n]).abs().max()) / sc)
        e_fp32.append(float((att[0][b, :, :n].double() - r_exact[b, :, :n]).abs().max()) / sc)
    rel_z = float((z[1] - z[0]).norm() / z[0].norm())
    record_parity({"test": "tensor_core_attention", "B": B, "T": T, "lens": lens, "vs_fp64_on_fp16_inputs": e_kernel,
                   "vs_fp32_kernel": e_round, "fp32_kernel_vs_fp64": e_fp32, "z_rel_l2_vs_fp32_attention": rel_z})
    assert max(e_fp32) <= 2e-4, e_fp32
    assert max(e_kernel) <= 5e-4, e_kernel
    assert max(e_round) <= 8e-2, e_round
    assert rel_z <= 1e-2, rel_z


@pytest.mark.gpu
def test_front_end_graph_replay_is_bit_identical_to_direct_launches():
    """rvcsyn_set_option("graphs"): small batches copy their inputs to fixed workspace places and replay the front end's launch
    sequence as a CUDA graph.  Capture call, replay call with OTHER inputs in other tensors, and direct launches agree bit
    for bit, in fp32 and fast mode; a changed option invalidates the graph."""
    import rvc_b200

    dev = torch.device("cuda", 0)
    torch.manual_seed(80)
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval().to(dev)
    B, T = 2, 150
    h = m._ensure_handle(dev)

    def inputs(seed):
        gen = torch.Generator().manual_seed(seed)
        return dict(phone=torch.randn(B, T, 768, generator=gen).to(dev), lens=torch.tensor([150, 77]).to(dev),
                    pitch=torch.randint(1, 256, (B, T), generator=gen).to(dev), f0=O.synth_f0(B, T, gen).to(dev),
                    sid=torch.tensor([3, 100]).to(dev), nzp=torch.randn(B, 192, T, generator=gen).to(dev),
                    nsrc=torch.randn(B, T * 480, generator=gen).to(dev))

    def run(i, mode):
        o, _, (z, _, _, _) = m.infer(i["phone"], i["lens"], i["pitch"], i["f0"], i["sid"], noise_zp=i["nzp"], noise_src=i["nsrc"], mode=mode)
        return o.clone(), z.clone()

    a, b = inputs(1), inputs(2)
    for mode in ("fp32", "fast"):
        h.set_option("graphs", 1)
        oa1, za1 = run(a, mode)   # capture
        ob1, zb1 = run(b, mode)   # replay on other inputs
        oa2, za2 = run(a, mode)
        n_graph = m.last_launch_count
        h.set_option("graphs", 0)
        oa0, za0 = run(a, mode)
        ob0, zb0 = run(b, mode)
        n_direct = m.last_launch_count
        assert torch.equal(za1, za0) and torch.equal(za2, za0) and torch.equal(zb1, zb0)
        assert torch.equal(oa1, oa0) and torch.equal(ob1, ob0)
        assert not torch.equal(za0, zb0)
        assert n_graph == n_direct + 1, (n_graph, n_direct)
    h.set_option("graphs", 1)
    _, z1 = run(a, "fast")
    h.set_option("att_mma", 0)  # another kernel sequence: the cached graph must not be replayed
    _, z2 = run(a, "fast")
    h.set_option("att_mma", 1)
    _, z3 = run(a, "fast")
    assert torch.equal(z1, z3) and not torch.equal(z1, z2)
