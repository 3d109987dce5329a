"""b2_paged_decode_append (include/b200attn.h): the decode step's KV write and its attention in one call, one launch for
the single-pass grid.  Bar: `out` AND both caches bit-identical to b2_reshape_and_cache followed by b2_paged_decode, and
`out` inside the decode bound against the CPU oracle.

The cache is filled with every token but the newest of each row; the newest token's slot is NaN-poisoned (bf16) / holds
the e4m3 NaN byte (fp8), so an append that lands after -- or is invisible to -- the CTA's own TMA load of that tile
turns the row into NaN.  Lengths of 1..65 put the new token into the FIRST tile a CTA loads (the copy is issued right
behind the store: the tightest ordering case), lengths of block_size * n + 1 put it into a fresh block."""
import pytest
import torch

import b200attn
import paged_attention_ref as R

pytestmark = pytest.mark.gpu

ATOL, RTOL = 2e-3, 1e-2


@pytest.fixture(autouse=True, params=[1, 2], ids=["producer-store", "consumer-store"])
def append_mode(request, monkeypatch):
    """Both ways the fused path can store the row (B2_APPEND_MODE, csrc/paged_decode.cu): 1 = the producer warp in front of
    its last stage (default), 2 = consumer warp 0 at CTA start with an mbarrier hand-over to the producer."""
    monkeypatch.setenv("B2_APPEND_MODE", str(request.param))
    return request.param


def _case(dev, lens, Hq, Hkv, bs, fp8, seed, pad_rows=()):
    d = 128
    B = len(lens)
    ctx = max(max(lens), 1)
    g = torch.Generator().manual_seed(seed)
    max_nb = (ctx + bs - 1) // bs + 1
    num_blocks = B * max_nb + 9
    perm = torch.randperm(num_blocks, generator=g).to(torch.int32)
    bt = perm[:B * max_nb].reshape(B, max_nb).contiguous()
    seq_lens = torch.tensor(lens, dtype=torch.int32)
    q = torch.randn(B, Hq, d, generator=g).to(torch.bfloat16)
    shape = (num_blocks, Hkv, bs, d)
    if fp8:
        ks, vs = 5.5 / 448.0, 6.5 / 448.0
        kc_ref = torch.full(shape, 0x7F, dtype=torch.uint8)          # 0x7f: the e4m3 NaN
        vc_ref = torch.full(shape, 0x7F, dtype=torch.uint8)
        scales = (torch.tensor([ks], device=dev), torch.tensor([vs], device=dev))
        okw = dict(kv_dtype=R.KV_FP8_E4M3, k_scale=ks, v_scale=vs)
    else:
        kc_ref = torch.full(shape, float("nan"), dtype=torch.bfloat16)
        vc_ref = torch.full(shape, float("nan"), dtype=torch.bfloat16)
        scales = (None, None)
        okw = {}
    k_new = torch.zeros(B, Hkv, d, dtype=torch.bfloat16)
    v_new = torch.zeros(B, Hkv, d, dtype=torch.bfloat16)
    slot_new = torch.full((B,), -1, dtype=torch.int64)
    for b, L in enumerate(lens):
        if L == 0:
            continue
        j = torch.arange(L)
        slots = bt[b, j // bs].long() * bs + (j % bs)
        k = torch.randn(L, Hkv, d, generator=g).to(torch.bfloat16)
        v = torch.randn(L, Hkv, d, generator=g).to(torch.bfloat16)
        old = L if b in pad_rows else L - 1                         # a padding row's newest token is already cached
        if old:
            if fp8:
                R.reshape_and_cache_ref(k[:old], v[:old], kc_ref, vc_ref, slots[:old], R.KV_FP8_E4M3, ks, vs)
            else:
                R.reshape_and_cache_ref(k[:old], v[:old], kc_ref, vc_ref, slots[:old])
        k_new[b], v_new[b] = k[L - 1], v[L - 1]
        if b not in pad_rows:
            slot_new[b] = slots[L - 1]
    return dict(q=q.to(dev), k_new=k_new.to(dev), v_new=v_new.to(dev), slot_new=slot_new.to(dev), bt=bt.to(dev),
                seq_lens=seq_lens.to(dev), kc0=kc_ref.to(dev), vc0=vc_ref.to(dev), scales=scales, okw=okw,
                host=dict(q=q, k_new=k_new, v_new=v_new, slot_new=slot_new, bt=bt, seq_lens=seq_lens, kc=kc_ref, vc=vc_ref,
                          fp8=fp8, ks=okw.get("k_scale"), vs=okw.get("v_scale")))


def _bits(t):
    return t.contiguous().view(torch.uint8).cpu()


def _check(c, splits, oracle=True):
    ks, vs = c["scales"]
    # the two calls the append stands for
    kc_a, vc_a = c["kc0"].clone(), c["vc0"].clone()
    b200attn.reshape_and_cache(c["k_new"], c["v_new"], kc_a, vc_a, c["slot_new"], ks, vs)
    out_a = b200attn.paged_decode(c["q"], kc_a, vc_a, c["bt"], c["seq_lens"], k_scale=ks, v_scale=vs, num_splits=splits)
    for rep in range(2):                                            # twice: same bits on fresh caches
        kc_b, vc_b = c["kc0"].clone(), c["vc0"].clone()
        out_b = torch.full_like(out_a, float("nan"))
        b200attn.paged_decode_append(c["q"], c["k_new"], c["v_new"], c["slot_new"], kc_b, vc_b, c["bt"], c["seq_lens"],
                                     k_scale=ks, v_scale=vs, out=out_b, num_splits=splits)
        torch.cuda.synchronize()
        assert torch.equal(_bits(kc_b), _bits(kc_a)) and torch.equal(_bits(vc_b), _bits(vc_a)), "caches differ"
        assert torch.isfinite(out_b.float()).all(), "non-finite output: the appended row was not seen by the decode"
        assert torch.equal(_bits(out_b), _bits(out_a)), f"out differs from the two-call result (rep {rep})"
    if oracle:
        h = c["host"]
        kc, vc = h["kc"].clone(), h["vc"].clone()
        if h["fp8"]:
            R.reshape_and_cache_ref(h["k_new"], h["v_new"], kc, vc, h["slot_new"], R.KV_FP8_E4M3, h["ks"], h["vs"])
        else:
            R.reshape_and_cache_ref(h["k_new"], h["v_new"], kc, vc, h["slot_new"])
        assert torch.equal(_bits(kc_b), _bits(kc)) and torch.equal(_bits(vc_b), _bits(vc)), "caches differ from the oracle's"
        ref = R.paged_decode_ref(h["q"], kc, vc, h["bt"], h["seq_lens"], return_f32=True, **c["okw"])
        err = (out_b.float().cpu() - ref).abs()
        assert bool((err <= ATOL + RTOL * ref.abs()).all()), f"max err {float(err.max()):.3e}"


EDGE_LENS = [1, 2, 17, 63, 64, 65, 129, 300, 16, 33, 0, 257]


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("Hq,Hkv", [(8, 2), (32, 8), (16, 1), (4, 4)])
def test_append_fused_single_pass(cuda, Hq, Hkv, fp8):
    assert b200attn.append_is_fused(Hq, Hkv, 1)
    _check(_case(cuda, EDGE_LENS, Hq, Hkv, 16, fp8, seed=Hq + Hkv), splits=1)


@pytest.mark.parametrize("bs", [8, 32])
@pytest.mark.parametrize("fp8", [False, True])
def test_append_fused_block_sizes(cuda, bs, fp8):
    lens = [1, bs, bs + 1, 2 * bs + 1, 3 * bs - 1, 70, 200]
    _check(_case(cuda, lens, 8, 2, bs, fp8, seed=bs), splits=1)


@pytest.mark.parametrize("splits,Hq,Hkv", [(0, 8, 2), (3, 8, 2), (1, 40, 2), (0, 32, 8)])
def test_append_unfused_schedules(cuda, splits, Hq, Hkv):
    """Schedules that cut a row over several CTAs (dynamic, split-K) and G > 16 (two head-group CTAs per kv head):
    the call launches the KV write in front of the decode grid -- same results."""
    assert not b200attn.append_is_fused(Hq, Hkv, splits)
    _check(_case(cuda, EDGE_LENS, Hq, Hkv, 16, False, seed=splits + Hq), splits=splits)


def test_append_padding_rows_and_strides(cuda):
    """slot < 0 rows write nothing (their newest token is already cached here); k_new / v_new as strided views of a fused
    q|k|v buffer, the way an engine holds them."""
    c = _case(cuda, [40, 5, 77, 130, 64], 8, 2, 16, False, seed=11, pad_rows=(1, 3))
    B, Hq, Hkv, d = 5, 8, 2, 128
    qkv = torch.zeros(B, Hq + 2 * Hkv, d, dtype=torch.bfloat16, device=cuda)
    qkv[:, :Hq], qkv[:, Hq:Hq + Hkv], qkv[:, Hq + Hkv:] = c["q"], c["k_new"], c["v_new"]
    c["q"], c["k_new"], c["v_new"] = qkv[:, :Hq], qkv[:, Hq:Hq + Hkv], qkv[:, Hq + Hkv:]
    _check(c, splits=1)


def test_append_many_short_rows_config_shape(cuda):
    """256 rows x 8 kv heads of 1..130 tokens: 2048 CTAs whose first (often only) tile holds the appended row."""
    g = torch.Generator().manual_seed(5)
    lens = torch.randint(1, 131, (256,), generator=g).tolist()
    _check(_case(cuda, lens, 32, 8, 16, False, seed=6), splits=1, oracle=False)
    _check(_case(cuda, lens[:128], 32, 8, 16, True, seed=7), splits=1, oracle=False)


@pytest.mark.parametrize("fp8", [False, True])
@pytest.mark.parametrize("splits", [1, 0])
def test_plan_run_append_resident_tables(cuda, fp8, splits):
    """b2_decode_plan_run_append over device-RESIDENT block tables: rows whose new token opens a block (length = n * bs + 1)
    arrive with a stale table entry for it (a wrong block id here), exactly what an engine's resident table holds before
    the step.  Bar: out, both caches and the table bit-identical to b2_reshape_and_cache_decode + b2_decode_plan_run."""
    from b200attn import ops
    bs = 16
    lens = [1, 17, 33, 16, 65, 64, 129, 300, 2, 49]
    c = _case(cuda, lens, 8, 2, bs, fp8, seed=21 + splits)
    ks, vs = c["scales"]
    stale = c["bt"].clone()
    for b, L in enumerate(lens):
        if (L - 1) % bs == 0:
            stale[b, (L - 1) // bs] = 3                       # not the block the slot names
    res = []
    for fused in (False, True):
        kc, vc, bt = c["kc0"].clone(), c["vc0"].clone(), stale.clone()
        plan = ops.DecodePlan(kc, vc, len(lens), 8, bt.shape[1], k_scale=ks, v_scale=vs)
        out = torch.full((len(lens), 8, 128), float("nan"), dtype=torch.bfloat16, device=cuda)
        if fused:
            plan.run_append(c["q"], c["k_new"], c["v_new"], c["slot_new"], bt, c["seq_lens"], out, splits)
        else:
            ops.reshape_and_cache_decode(c["k_new"], c["v_new"], kc, vc, c["slot_new"], bt, c["seq_lens"], ks, vs)
            plan.run(c["q"], bt, c["seq_lens"], out, splits)
        torch.cuda.synchronize()
        res.append((out, kc, vc, bt))
        plan.close()
    (o0, k0, v0, t0), (o1, k1, v1, t1) = res
    assert torch.equal(t0.cpu(), c["bt"].cpu()) and torch.equal(t1.cpu(), c["bt"].cpu()), "tables not brought up to date"
    assert torch.equal(_bits(k1), _bits(k0)) and torch.equal(_bits(v1), _bits(v0)), "caches differ"
    assert torch.isfinite(o1.float()).all()
    assert torch.equal(_bits(o1), _bits(o0)), "out differs from the two-call result"


def test_append_refused_call_writes_nothing(cuda):
    """A call that fails its argument checks (workspace too small for a split-K schedule) must not have launched the
    KV write either: the caches keep their bytes."""
    c = _case(cuda, [300, 200, 129], 8, 2, 16, False, seed=31)
    kc, vc = c["kc0"].clone(), c["vc0"].clone()
    small = torch.zeros(64, dtype=torch.uint8, device=cuda)
    with pytest.raises(ValueError):
        b200attn.paged_decode_append(c["q"], c["k_new"], c["v_new"], c["slot_new"], kc, vc, c["bt"], c["seq_lens"],
                                     workspace=small, num_splits=3)
    # through the C ABI (the Python wrapper sizes workspaces itself): the library's own refusal
    import ctypes as C
    from b200attn import _lib
    lib = _lib.load()
    q, bt = c["q"], c["bt"]
    rc = lib.b2_paged_decode_append(
        C.c_void_p(torch.empty_like(q).data_ptr()), C.c_void_p(q.data_ptr()), C.c_void_p(c["k_new"].data_ptr()),
        C.c_void_p(c["v_new"].data_ptr()), C.c_void_p(c["slot_new"].data_ptr()), 2 * 128, 2 * 128,
        C.c_void_p(kc.data_ptr()), C.c_void_p(vc.data_ptr()), C.c_void_p(bt.data_ptr()), C.c_void_p(c["seq_lens"].data_ptr()),
        3, 8, 2, 128, 16, bt.shape[1], kc.shape[0], 8 * 128, 8 * 128, kc.stride(0), kc.stride(1), 128 ** -0.5, 0, None, None,
        C.c_void_p(small.data_ptr()), 64, 3, C.c_void_p(torch.cuda.current_stream(cuda).cuda_stream))
    assert rc == _lib.B2_ENOMEM, rc
    torch.cuda.synchronize()
    assert torch.equal(_bits(kc), _bits(c["kc0"])) and torch.equal(_bits(vc), _bits(c["vc0"])), "a refused call wrote the cache"
