"""GPU: the tensor-parallel exchange kernels of csrc/comm.cu (`--tensor-parallel-size N`, reference
benchmark-vllm/models.py:11) on ONE GPU.

* tp = 1: b2_comm_allreduce_rms_norm must equal b2_rms_norm with the fused residual add (residual bit for bit), and
  b2_comm_allgather_cols must be a copy.
* tp = 2 emulated in one process: two regions on the same device, "rank 0" and "rank 1" launched on two streams, so
  the flag protocol, the epoch counter, the slot alternation and the peer loads / stores run for real (the peers
  just happen to share a device).  Checked against a plain fp32 torch sum.  It runs in a SUBPROCESS: a protocol bug
  traps the kernel after 10 s and poisons the CUDA context, which must not take the rest of the suite with it.
The two-process CUDA-IPC path over NVLink is exercised by scripts/tp_check.py / bench.py --gpus N (needs >= 2 GPUs)."""
import ctypes as C
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


class Regions:
    """n comm regions on the current device + one b2_comm handle per emulated rank."""

    def __init__(self, n, max_rows, H, grows, gcols):
        from b200attn import _lib
        from b200attn.tp import _device_view
        self.lib, self.n, self.H, self.max_rows = _lib.load(), n, H, max_rows
        lib = self.lib
        nbytes = int(lib.b2_comm_region_bytes(n, max_rows, H, grows, gcols))
        self.ptrs = []
        for _ in range(n):
            p, h = C.c_void_p(), C.create_string_buffer(64)
            _lib.check(lib.b2_comm_alloc(nbytes, C.byref(p), h))
            self.ptrs.append(p.value)
        arr = (C.c_void_p * n)(*self.ptrs)
        self.comms = []
        for r in range(n):
            c = lib.b2_comm_create(r, n, arr, nbytes, max_rows, H, grows, gcols)
            assert c, lib.b2_last_error()
            self.comms.append(c)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.inputs = [[_device_view(lib.b2_comm_input(c, s), max_rows * H * 2, dev).view(torch.bfloat16).view(max_rows, H)
                        for s in (0, 1)] for c in self.comms]
        self.gathered = [_device_view(lib.b2_comm_gathered(c), grows * gcols * 2, dev).view(torch.bfloat16).view(grows, gcols)
                         if grows else None for c in self.comms]

    def allreduce(self, r, slot, out, resid, w, T, eps, stream):
        from b200attn import _lib
        _lib.check(self.lib.b2_comm_allreduce_rms_norm(self.comms[r], slot, out.data_ptr(),
                                                       None if resid is None else resid.data_ptr(),
                                                       None if w is None else w.data_ptr(), T, self.H, self.H, eps,
                                                       stream.cuda_stream))

    def push(self, r, part, out, resid, w, eps, stream):
        from b200attn import _lib
        T = part.shape[0]
        _lib.check(self.lib.b2_comm_push_allreduce_rms_norm(self.comms[r], part.data_ptr(), part.stride(0), out.data_ptr(),
                                                            None if resid is None else resid.data_ptr(),
                                                            None if w is None else w.data_ptr(), T, self.H, self.H, eps,
                                                            stream.cuda_stream))

    def allgather(self, r, local, stream):
        from b200attn import _lib
        T, cols = local.shape
        _lib.check(self.lib.b2_comm_allgather_cols(self.comms[r], local.data_ptr(), local.stride(0), T, cols,
                                                   stream.cuda_stream))

    def close(self):
        torch.cuda.synchronize()
        for c in self.comms:
            self.lib.b2_comm_destroy(c)
        for p in self.ptrs:
            self.lib.b2_comm_free(p)


@pytest.mark.parametrize("H", [512, 4096, 8192])
def test_comm_tp1_equals_rms_norm(cuda, H):
    from b200attn import ops
    T = 37
    R = Regions(1, 64, H, 64, 1024)
    try:
        st = torch.cuda.current_stream()
        x = torch.randn(T, H, device="cuda").to(torch.bfloat16)
        w = torch.randn(H, device="cuda").to(torch.bfloat16)
        resid = torch.randn(T, H, device="cuda").to(torch.bfloat16)
        for slot in (0, 1, 0):
            R.inputs[0][slot][:T].copy_(x)
            r1, r2 = resid.clone(), resid.clone()
            out = torch.empty_like(x)
            R.allreduce(0, slot, out, r1, w, T, 1e-5, st)
            ref = ops.rms_norm(x, w, 1e-5, residual=r2)
            assert torch.equal(r1.view(torch.int16), r2.view(torch.int16))           # the residual stream: bit-exact
            # the norm: the two kernels sum the squares in a different thread order -> at most one bf16 ulp apart
            assert bool(((out.float() - ref.float()).abs() <= 2.0 ** -7 * ref.float().abs()).all())
            assert (out.view(torch.int16) != ref.view(torch.int16)).float().mean() < 0.01
            r3, out3 = resid.clone(), torch.empty_like(x)
            R.push(0, x, out3, r3, w, 1e-5, st)                                      # push variant: same arithmetic
            assert torch.equal(out3.view(torch.int16), out.view(torch.int16))
            assert torch.equal(r3.view(torch.int16), r2.view(torch.int16))
        # no weight: plain sum (+ residual); no residual: plain norm
        out = torch.empty_like(x)
        R.allreduce(0, 1, out, None, None, T, 1e-5, st)
        assert torch.equal(out.view(torch.int16), x.view(torch.int16))
        loc = torch.randn(T, 1024, device="cuda").to(torch.bfloat16)
        R.allgather(0, loc, st)
        assert torch.equal(R.gathered[0][:T].view(torch.int16), loc.view(torch.int16))
    finally:
        R.close()


def _two_rank_protocol():
    """Body of the subprocess: 2 emulated ranks on one device, 40 alternating exchanges + gathers."""
    sys.path.insert(0, str(ROOT / "vllm-backend-public_b200"))
    torch.manual_seed(0)
    H, T, V = 1024, 8, 512
    R = Regions(2, 16, H, 16, V)
    s = [torch.cuda.Stream(), torch.cuda.Stream()]
    w = torch.randn(H, device="cuda").to(torch.bfloat16)
    worst = 0.0
    for it in range(60):
        slot = (it // 3) & 1
        parts = [torch.randn(T, H, device="cuda").to(torch.bfloat16) for _ in range(2)]
        resid0 = torch.randn(T, H, device="cuda").to(torch.bfloat16)
        torch.cuda.synchronize()
        outs, resids = [], []
        if it % 7 == 3:
            parts[0][2, 5:9] = -0.0                      # the sentinel value itself must survive the push protocol
        for r in range(2):
            with torch.cuda.stream(s[r]):
                out, rs = torch.empty(T, H, dtype=torch.bfloat16, device="cuda"), resid0.clone()
                if it % 3 == 2:                          # every third exchange through the pull variant
                    R.inputs[r][slot][:T].copy_(parts[r])
                    R.allreduce(r, slot, out, rs, w, T, 1e-5, s[r])
                else:
                    R.push(r, parts[r], out, rs, w, 1e-5, s[r])
                outs.append(out)
                resids.append(rs)
        torch.cuda.synchronize()
        h = (parts[0].float() + parts[1].float() + resid0.float()).to(torch.bfloat16)
        hf = h.float()
        ref = (w.float() * (hf * torch.rsqrt((hf * hf).mean(dim=1, keepdim=True) + 1e-5))).to(torch.bfloat16)
        for r in range(2):
            assert torch.equal(resids[r].view(torch.int16), h.view(torch.int16)), f"iteration {it}: residual of rank {r}"
            worst = max(worst, float((outs[r].float() - ref.float()).abs().max()))
        assert torch.equal(outs[0].view(torch.int16), outs[1].view(torch.int16)), "ranks must compute the same bits"
        if it % 5 == 4:                                  # a gather between exchanges, as after the lm_head
            loc = [torch.randn(T, V // 2, device="cuda").to(torch.bfloat16) for _ in range(2)]
            torch.cuda.synchronize()
            for r in range(2):
                with torch.cuda.stream(s[r]):
                    R.allgather(r, loc[r], s[r])
            torch.cuda.synchronize()
            full = torch.cat(loc, dim=1)
            for r in range(2):
                assert torch.equal(R.gathered[r][:T].view(torch.int16), full.view(torch.int16)), f"gather on rank {r}"
    assert worst <= 0.07, worst            # one bf16 ulp at |x| <= 8
    R.close()
    print("two-rank protocol ok, worst", worst)


def test_comm_two_emulated_ranks_on_one_gpu(cuda):
    r = subprocess.run([sys.executable, str(Path(__file__).resolve()), "two-rank"], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0 and "two-rank protocol ok" in r.stdout, (r.stdout[-2000:], r.stderr[-2000:])


if __name__ == "__main__" and sys.argv[1:] == ["two-rank"]:
    _two_rank_protocol()
