"""CPU tests of the host side: C-ABI exports, weight folding/packing images, polyphase
decomposition, chunk geometry, sharding (incl. a world_size-2 gloo run).  No kernel is launched."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import nsf_hifigan_oracle as O
from helpers import build_module, golden_state, load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_exports_every_declared_symbol(built_lib):
    from rvc_b200 import _lib

    declared = set()
    for header, prefix in (("rvcdec.h", "rvcdec"), ("rvcsyn.h", "rvcsyn"), ("rvcrefine.h", "rvcrefine"), ("rvcrmvpe.h", "rvcrmvpe"), ("rvchubert.h", "rvchubert")):
        hdr = open(os.path.join(ROOT, "include", header)).read()
        found = set(re.findall(r"\b(%s_[a-z_0-9]+)\s*\(" % prefix, hdr))
        assert found, f"no declarations parsed in {header}"
        declared |= found
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for sym in declared:
        assert getattr(built_lib, sym) is not None
    assert b"sm_100a" in built_lib.rvcdec_version()


def test_module_state_dict_matches_reference_naming():
    d, cfg = load_golden("v2_48k_B2_T24")
    m = build_module(cfg, int(d["init_seed"]))
    sd = m.state_dict()
    assert len(sd) == 243  # 2 + 2 + 4*3 + 4*2 + 12*6*3 + 1 + 2
    assert {k[4:] for k in d.files if k.startswith("ck::")} == set(sd.keys())
    assert m.upp == 480 and m.num_kernels == 3 and m.num_upsamples == 4
    assert len(m.ups) == 4 and len(m.noise_convs) == 4 and len(m.resblocks) == 12
    # legacy checkpoint naming (weight_g / weight_v) loads, as rvc/infer/infer.py:485 relies on
    legacy = {k.replace(".parametrizations.weight.original0", ".weight_g").replace(".parametrizations.weight.original1", ".weight_v"): v for k, v in sd.items()}
    m2 = build_module(cfg, 1)
    res = m2.load_state_dict(legacy, strict=False)
    assert not res.missing_keys and not res.unexpected_keys
    assert torch.equal(m2.ups[0].weight, m.ups[0].weight)
    m2.remove_weight_norm()  # must not raise (the reference's does)
    assert "ups.0.weight" in m2.state_dict()
    assert torch.allclose(m2.ups[0].weight, m.ups[0].weight, atol=1e-7)


def test_forward_without_cuda_fails_loudly():
    d, cfg = load_golden("tiny_24k_B2_T40")
    m = build_module(cfg, 0)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.zeros(1, cfg.initial_channel, 4), torch.zeros(1, 4))


def _conv_from_image(x, wf, shifts, up, cout):
    """Evaluate a packed layer image exactly as the kernels do: D[t,n] = sum_tap x[t+shift] @ wf[tap]."""
    B, Cin, L = x.shape
    halo = int(max(abs(int(s)) for s in shifts))
    xp = F.pad(x, (halo, halo)).double()
    N = wf.shape[2]
    D = torch.zeros(B, L, N, dtype=torch.float64)
    for tap, s in enumerate(shifts):
        D += xp[:, :, halo + int(s): halo + int(s) + L].transpose(1, 2) @ torch.from_numpy(wf[tap]).double()
    return D.reshape(B, L, up, cout).permute(0, 3, 1, 2).reshape(B, cout, L * up)


def _decode16(img, fmt):
    if fmt == "f16":
        return torch.from_numpy(img.view(np.float16).astype(np.float32))
    return (torch.from_numpy(img.astype(np.int32)) << 16).view(torch.float32)


@pytest.mark.parametrize("name", ["tiny_24k_B2_T40", "v2_40k_B1_T32"])
def test_packed_images_reproduce_the_layers(built_lib, name):
    from rvc_b200 import _lib

    d, cfg = load_golden(name)
    sd = golden_state(d, cfg)
    w = O.folded_weights(sd)
    h = _lib.Handle(_lib.make_config(*cfg.ctor_args(), cfg.gin_channels, cfg.sr), 0)
    h.pack_host(sd)
    gen = torch.Generator().manual_seed(0)
    layers = ["conv_pre", "ups.0", "ups.1", "ups.2", "ups.3", "resblocks.0.convs1.2", "resblocks.4.convs1.1",
              "resblocks.8.convs2.0", "resblocks.11.convs1.2"]
    for layer in layers:
        meta = h.packed(layer, 5)
        Cin, N, cout, up, ntaps, halo, n_tile, kc = (int(v) for v in meta)
        wf = h.packed(layer, 0).reshape(ntaps, Cin, N)
        shifts = h.packed(layer, 2)
        bias = torch.from_numpy(h.packed(layer, 6)).double()
        x = torch.randn(2, Cin, 37, generator=gen)
        got = _conv_from_image(x, wf, shifts, up, cout) + bias.reshape(up, cout)[0].reshape(1, -1, 1)
        if layer == "conv_pre":
            exp = F.conv1d(x.double(), w["conv_pre.weight"].double(), w["conv_pre.bias"].double(), padding=3)
        elif layer.startswith("ups."):
            i = int(layer[4:])
            u, k = cfg.upsample_rates[i], cfg.upsample_kernel_sizes[i]
            exp = F.conv_transpose1d(x.double(), w[layer + ".weight"].double(), w[layer + ".bias"].double(), stride=u, padding=(k - u) // 2)
        else:
            j = int(layer.split(".")[1]) % 3
            m = int(layer.split(".")[3])
            k = cfg.resblock_kernel_sizes[j]
            dil = cfg.resblock_dilation_sizes[j][m] if ".convs1." in layer else 1
            exp = F.conv1d(x.double(), w[layer + ".weight"].double(), w[layer + ".bias"].double(), dilation=dil, padding=O.get_padding(k, dil))
        assert got.shape == exp.shape, layer
        assert float((got - exp).abs().max()) < 2e-6, layer  # fp32 weight-norm fold rounding only
        # tensor-core image: [n_tile][tap][kchunk][slice][n][8], equal to the fp32 image rounded to 16 bits
        if n_tile:
            for kind, fmt in ((1, "f16"), (4, "bf16")):
                img = _decode16(h.packed(layer, kind), fmt).reshape(N // n_tile, ntaps, Cin // kc, kc // 8, n_tile, 8)
                back = img.permute(1, 2, 3, 5, 0, 4).reshape(ntaps, Cin, N)
                ref16 = torch.from_numpy(wf)
                ref16 = ref16.half().float() if fmt == "f16" else ref16.bfloat16().float()
                assert torch.equal(back, ref16), (layer, fmt)
            masks = h.packed(layer, 3)
            for t in range(N // n_tile):
                for tap in range(ntaps):
                    nz = bool(np.any(wf[tap][:, t * n_tile:(t + 1) * n_tile] != 0))
                    assert bool((int(masks[t]) >> tap) & 1) == nz, (layer, t, tap)
        elif name.startswith("v2_"):
            pytest.fail(f"{layer}: v2 layer without a tensor-core tiling")
    h.close()


def test_v2_48k_tilings(built_lib):
    """Every conv of the 48 kHz v2 decoder gets a tcgen05 tiling; ups0 skips its all-zero tap blocks."""
    from rvc_b200 import _lib

    d, cfg = load_golden("v2_48k_B2_T24")
    sd = golden_state(d, cfg)
    h = _lib.Handle(_lib.make_config(*cfg.ctor_args(), cfg.gin_channels, cfg.sr), 0)
    h.pack_host(sd)
    # default: 128-wide N tiles (accumulators double-buffered in TMEM)
    expect = {"conv_pre": (192, 512, 128, 64), "ups.0": (512, 3072, 128, 64), "ups.1": (256, 1280, 128, 64),
              "ups.2": (128, 128, 128, 64), "ups.3": (64, 64, 64, 64), "resblocks.0.convs1.0": (256, 256, 128, 64),
              "resblocks.3.convs1.0": (128, 128, 128, 64), "resblocks.6.convs2.1": (64, 64, 64, 64),
              "resblocks.9.convs1.2": (32, 32, 32, 32)}
    for layer, (Cin, N, nt, kc) in expect.items():
        meta = h.packed(layer, 5)
        assert (int(meta[0]), int(meta[1]), int(meta[6]), int(meta[7])) == (Cin, N, nt, kc), layer
    m0 = h.packed("ups.0", 3)
    assert len(m0) == 24 and all(bin(int(v)).count("1") == 2 for v in m0)  # exactly 2 of 3 shifts per phase half
    # option max_ntile = 256 (takes effect at the next pack): the widest tiles the kernels support
    h.set_option("max_ntile", 256)
    h.pack_host(sd)
    for layer, (Cin, N, nt, kc) in {"conv_pre": (192, 512, 256, 64), "ups.0": (512, 3072, 256, 64),
                                     "resblocks.0.convs1.0": (256, 256, 256, 64), "resblocks.3.convs1.0": (128, 128, 128, 64)}.items():
        meta = h.packed(layer, 5)
        assert (int(meta[0]), int(meta[1]), int(meta[6]), int(meta[7])) == (Cin, N, nt, kc), layer
    m0 = h.packed("ups.0", 3)
    assert len(m0) == 12 and all(bin(int(v)).count("1") == 2 for v in m0)
    h.close()


def test_phase_major_upsample_images(built_lib):
    """The rate-10 / 12 upsamples also get a phase-major tensor-core image: tile (co/16, half of the phases), columns
    (slice, phase, channel); it is the standard polyphase image with permuted columns, and every tile still skips one of
    the three input shifts."""
    from rvc_b200 import _lib

    d, cfg = load_golden("v2_48k_B2_T24")
    sd = golden_state(d, cfg)
    h = _lib.Handle(_lib.make_config(*cfg.ctor_args(), cfg.gin_channels, cfg.sr), 0)
    h.pack_host(sd)
    for i, (u, C) in enumerate([(12, 256), (10, 128)]):
        base, pm = f"ups.{i}", f"ups.{i}(phase-major)"
        Cin, N, cout, up, ntaps, halo, n_tile, kc = (int(v) for v in h.packed(pm, 5))
        PH = u // 2
        assert (N, cout, up, n_tile, kc) == (u * C, C, u, PH * 16, 64)
        wf = torch.from_numpy(h.packed(base, 0).reshape(ntaps, Cin, N))
        r = torch.arange(u).repeat_interleave(C)
        co = torch.arange(C).repeat(u)
        n_new = ((co // 16) * 2 + r // PH) * n_tile + (((co // 8) % 2) * PH + r % PH) * 8 + co % 8
        assert sorted(n_new.tolist()) == list(range(N))
        img = _decode16(h.packed(pm, 4), "bf16").reshape(N // n_tile, ntaps, Cin // kc, kc // 8, n_tile, 8)
        back = img.permute(1, 2, 3, 5, 0, 4).reshape(ntaps, Cin, N)
        assert torch.equal(back[:, :, n_new], wf.bfloat16().float())
        bias = torch.from_numpy(h.packed(pm, 6))
        assert torch.equal(bias[n_new], torch.from_numpy(h.packed(base, 6)))
        masks = h.packed(pm, 3)
        assert len(masks) == N // n_tile and all(bin(int(v)).count("1") == 2 for v in masks)
    with pytest.raises(Exception):
        h.packed("ups.2(phase-major)", 5)
    h.close()


def test_chunk_plan_matches_reference_geometry():
    from rvc_b200 import chunking as ch

    # 5 min @ 100 frames/s with (x_pad,x_query,x_center,x_max) = (1,6,38,41): 7 split points, 8 chunks
    chunks = ch.plan_chunks(30000)
    assert len(chunks) == 8
    assert chunks[0].start == 0 and chunks[0].stop == 3800 + 200 and chunks[-1].stop == 30000 + 200
    covered = 0
    for c in chunks:
        assert c.trim == 100
        covered += c.frames - 2 * c.trim
    assert covered == 30000
    assert len(ch.plan_chunks(4100)) == 1 and len(ch.plan_chunks(4101)) == 2
    # trim + concat of per-chunk "decodes" of an identity signal reconstructs the signal
    sig = np.arange(30000 * 4, dtype=np.float32)
    padded = np.pad(sig, (400, 400), mode="reflect")
    dec = [padded[c.start * 4: c.stop * 4] for c in chunks]
    out = ch.stitch(dec, [c.trim * 4 for c in chunks])
    assert np.array_equal(out, sig)
    out_cf = ch.stitch(dec, [c.trim * 4 for c in chunks], crossfade=64)
    assert out_cf.shape == sig.shape and np.allclose(out_cf, sig, rtol=1e-6, atol=0.05)


def test_shard_units_balanced():
    from rvc_b200 import chunking as ch

    shards = ch.shard_units([3000] * 1024, 8)
    assert all(len(s) == 128 for s in shards)
    assert sorted(i for s in shards for i in s) == list(range(1024))
    shards = ch.shard_units([4000, 4000, 4000, 4000, 4000, 4000, 4000, 2200], 4)
    assert sorted(i for s in shards for i in s) == list(range(8))
    assert max(len(s) for s in shards) == 2


def test_gloo_world2_sharding(tmp_path):
    """The N>1 host path of bench.py (sharding + max-over-ranks reduction) on CPU with gloo."""
    script = tmp_path / "w.py"
    script.write_text(
        "import os, sys, torch, torch.distributed as dist\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "import bench\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "mine = bench.rank_units(10, r, w)\n"
        "t = bench.max_over_ranks(float(r + 1), torch.device('cpu'))\n"
        "tot = bench.sum_over_ranks(float(len(mine)), torch.device('cpu'))\n"
        "assert t == float(w) and tot == 10.0 * w and mine[0] == 10 * r, (t, tot, mine)\n"
        "print('ok', r, mine)\n"
    )
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", str(script)], capture_output=True, text=True, env=env, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert p.stdout.count("ok") == 2


def test_gloo_world2_long_form_decode_chunks_and_gather(tmp_path):
    """The torchrun form of BASELINE configs[4] on CPU with gloo, world size 2: chunking.decode_chunks(rank, world)
    deals the t_pad-overlap chunks longest-first, each rank decodes its share as one zero-padded batch through the
    host-buffer call (a stand-in model with the decoder's interface: sample s of a chunk depends only on the frames
    within +-3 of it, like a conv stack with a short receptive field), chunking.gather_chunks collects the results on
    rank 0 and the stitched utterance equals the single-rank run."""
    script = tmp_path / "lf.py"
    script.write_text(
        "import os, sys, numpy as np, torch, torch.distributed as dist\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "from rvc_b200 import chunking as ch\n"
        "class Fake:\n"
        "    upp = 4\n"
        "    def decode_host(self, x, f0, g, mode=None, device=None, out=None):\n"
        "        B, C, T = x.shape\n"
        "        k = torch.ones(1, C, 7) / (7 * C)\n"
        "        y = torch.nn.functional.conv1d(x, k, padding=3)          # (B, 1, T): receptive field +-3 frames\n"
        "        y = y.repeat_interleave(self.upp, dim=2) + 1e-3 * f0.repeat_interleave(self.upp, dim=1)[:, None]\n"
        "        return y + (0 if g is None else g.sum(dim=(1, 2))[:, None, None])\n"
        "dist.init_process_group('gloo')\n"
        "r, w = dist.get_rank(), dist.get_world_size()\n"
        "gen = torch.Generator().manual_seed(3)\n"
        "n_frames, pad = 900, 100\n"
        "x = torch.randn(1, 8, n_frames + 2 * pad, generator=gen); f0 = torch.rand(1, n_frames + 2 * pad, generator=gen) * 200\n"
        "g = torch.randn(1, 4, 1, generator=gen)\n"
        "chunks = ch.plan_chunks(n_frames, split_points=[250, 430, 700])\n"
        "mine = ch.decode_chunks(Fake(), x, f0, g, chunks, rank=r, world=w)\n"
        "assert set(mine) == set(ch.shard_units([c.frames for c in chunks], w)[r]), mine.keys()\n"
        "y = ch.gather_chunks(mine, chunks, Fake.upp)\n"
        "if r == 0:\n"
        "    single = ch.decode_chunks(Fake(), x, f0, g, chunks)\n"
        "    y1 = ch.stitch([single[c.index] for c in chunks], [c.trim * Fake.upp for c in chunks])\n"
        "    whole = Fake().decode_host(x, f0, g)[0, 0].numpy()[pad * 4: -pad * 4]\n"
        "    assert y.shape == (n_frames * 4,) and np.allclose(y, y1, atol=1e-6) and np.allclose(y, whole, atol=1e-6)\n"
        "else:\n"
        "    assert y is None\n"
        "print('ok', r, sorted(mine))\n"
    )
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29534", str(script)], capture_output=True, text=True, env=env, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert p.stdout.count("ok") == 2


def test_runner_deals_requeues_and_gathers_without_a_gpu():
    """runner.MultiGpuDecoder's host logic (LPT dealing, one pinned output buffer, re-queue of a failed worker's units)
    with stand-in handles: worker 1 fails its first batch, its units are re-dealt to worker 0 and every unit still
    arrives at its offset."""
    import numpy as np
    import torch
    from rvc_b200 import runner as R

    class FakeHandle:
        def __init__(self, dev):
            self.dev, self.calls = dev, 0

        def forward_host(self, x, f0, g, noise, seed, out, B, T, mode, cap):
            import ctypes as C
            self.calls += 1
            xs = np.ctypeslib.as_array(C.cast(x, C.POINTER(C.c_float)), shape=(B, 8, T))
            ys = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_float)), shape=(B, T * 4))
            ys[:] = np.repeat(xs.sum(1), 4, axis=1)

        def close(self):
            pass

    dec = R.MultiGpuDecoder.__new__(R.MultiGpuDecoder)
    dec.mode, dec.upp, dec.cin, dec.gin, dec.requeued = 1, 4, 8, 0, 0
    dec.workers = [R._Worker(0, FakeHandle(0)), R._Worker(1, FakeHandle(1))]
    dec.fail_hook = lambda dev, share: dev == 1
    x = torch.randn(1, 8, 1000)
    f0 = torch.zeros(1, 1000)
    units = [R.Unit(0, 0, 300), R.Unit(1, 250, 200), R.Unit(2, 400, 300), R.Unit(3, 650, 350)]
    out, offs = dec.decode_units(x, f0, None, units)
    for u in units:
        want = np.repeat(x[0, :, u.start:u.start + u.frames].sum(0).numpy(), 4)
        got = out[offs[u.index]: offs[u.index] + u.frames * 4].numpy()
        assert np.allclose(got, want, atol=1e-6), u
    st = dec.stats()
    assert st[1]["healthy"] is False and "injected" in st[1]["error"] and st[0]["units"] == 4 and dec.requeued == 2


def test_stride2_polyphase_images(built_lib):
    """The stride-2 upsamples also get a true polyphase image with the columns of a slice's two phases side by side
    (n = ((co/8)*2 + r)*8 + co%8): evaluated as the kernel does (input row t -> output rows 2t, 2t+1) it reproduces
    ConvTranspose1d(k=4, stride 2, padding 1) of hifigan_nsf.py:111-130."""
    from rvc_b200 import _lib

    d, cfg = load_golden("v2_48k_B2_T24")
    sd = golden_state(d, cfg)
    w = O.folded_weights(sd)
    h = _lib.Handle(_lib.make_config(*cfg.ctor_args(), cfg.gin_channels, cfg.sr), 0)
    h.pack_host(sd)
    gen = torch.Generator().manual_seed(1)
    for i, (Cin, C, nt) in ((2, (128, 64, 128)), (3, (64, 32, 64))):
        name = f"ups.{i}(polyphase-2)"
        meta = [int(v) for v in h.packed(name, 5)]
        assert meta[:5] == [Cin, 2 * C, C, 2, 3] and meta[6:] == [nt, 64], meta
        wf = h.packed(name, 0).reshape(3, Cin, 2 * C)
        shifts = [int(v) for v in h.packed(name, 2)]
        assert shifts == [-1, 0, 1]
        bias = torch.from_numpy(h.packed(name, 6)).double()
        x = torch.randn(2, Cin, 29, generator=gen)
        xp = F.pad(x, (1, 1)).double()
        D = torch.zeros(2, 29, 2 * C, dtype=torch.float64)
        for tap, s in enumerate(shifts):
            D += xp[:, :, 1 + s: 1 + s + 29].transpose(1, 2) @ torch.from_numpy(wf[tap]).double()
        D = D + bias
        got = D.reshape(2, 29, C // 8, 2, 8).permute(0, 2, 4, 1, 3).reshape(2, C, 58)  # (b, slice, c, t, r) -> (b, co, 2t + r)
        exp = F.conv_transpose1d(x.double(), w[f"ups.{i}.weight"].double(), w[f"ups.{i}.bias"].double(), stride=2, padding=1)
        assert float((got - exp).abs().max()) < 2e-6, name
        # every phase uses two of the three shifts: the shift -1 block is zero in the r = 1 columns and vice versa
        cols = np.arange(2 * C).reshape(C // 8, 2, 8)
        assert not np.any(wf[0][:, cols[:, 1].ravel()]) and not np.any(wf[2][:, cols[:, 0].ravel()])
    h.close()
