"""Row-sharded fit of ONE huge group across ranks (BASELINE config "1 B rows x 64 features on 8 GPUs").

Protocol (include/anofox_b200.h, "Row-sharded single group"); one process per GPU, torch.distributed is
the plumbing (NCCL on GPUs; the same code runs over gloo with a CPU backend in tests/):

  0. every rank finds the first valid row of its shard                 -> (found, shift candidate)
     all_gather; the candidate of the lowest rank that found one is the group's first valid row, the
     common origin every rank differences against (and the row of ols.rs:76-87's constant test)
  1. every rank accumulates the shifted moment block of its rows       -> moments [L], cdev [p]
  2. all_reduce(moments, SUM); all_reduce(cdev, MAX)                   (~35 KB at p = 64: latency-bound)
  3. every rank solves from the reduced block: identical inputs => identical results on every rank
  4. exact fits: when the first solve reports rss <= 1e-5 tss (device flag, no host read) every rank sums the
     squared residuals of its rows, all_reduce(SUM) of ONE double, and the group is solved again with it

The reference cannot run this case at all: its aggregate state would hold 520 GB of raw rows in one
std::vector.
"""
import ctypes as C

from . import _abi as abi


class CudaBackend:
    """Device compute of one rank through the C ABI (tensors are torch CUDA tensors)."""

    def __init__(self, n_rows_max, p, device):
        import torch
        from . import load_library
        self.torch, self.lib, self.p, self.device = torch, load_library(), p, device
        wb = self.lib.anofox_b200_workspace_bytes(1, n_rows_max, p)
        self.workspace = torch.empty(wb + 256, dtype=torch.uint8, device=device)
        self.ws_ptr, self.ws_bytes = (self.workspace.data_ptr() + 255) & ~255, wb
        self.L = int(self.lib.anofox_b200_moment_block_len(p))

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    @staticmethod
    def _p(t):
        return None if t is None else C.c_void_p(t.data_ptr())

    def _check(self, rc, err):
        if rc != 0:
            from . import B200Error
            raise B200Error(rc, err.message.decode(errors="replace"))

    def first_valid(self, y, x, w):
        torch = self.torch
        out = torch.empty(self.p + 2, dtype=torch.float64, device=self.device)  # [found, shift...]
        found = torch.empty(1, dtype=torch.int32, device=self.device)
        err = abi.AnofoxError()
        rc = self.lib.anofox_b200_first_valid_row_device(self._p(y), self._p(x), self._p(w), y.numel(), self.p,
                                                         C.c_void_p(out.data_ptr() + 8), self._p(found),
                                                         C.c_void_p(self.ws_ptr), self.ws_bytes, self._stream(),
                                                         C.byref(err))
        self._check(rc, err)
        out[0] = found[0].to(torch.float64)
        return out

    def moments(self, y, x, w, shift):
        torch = self.torch
        m = torch.empty(self.L, dtype=torch.float64, device=self.device)
        cd = torch.empty(self.p, dtype=torch.float64, device=self.device)
        err = abi.AnofoxError()
        rc = self.lib.anofox_b200_moments_device(self._p(y), self._p(x), self._p(w), y.numel(), self.p,
                                                 self._p(shift), self._p(m), self._p(cd), C.c_void_p(self.ws_ptr),
                                                 self.ws_bytes, self._stream(), C.byref(err))
        self._check(rc, err)
        return m, cd

    def solve(self, moments, cdev, shift, options, out_struct, flag=None, rss=None):
        """flag (int32[1]) given: the refining form -- first solve (rss None) or second solve with the exact RSS."""
        err = abi.AnofoxError()
        if flag is None:
            rc = self.lib.anofox_b200_solve_moments_device(self._p(moments), self._p(cdev), self._p(shift), self.p,
                                                           C.byref(options), C.byref(out_struct), self._stream(),
                                                           C.byref(err))
        else:
            rc = self.lib.anofox_b200_solve_moments_refine_device(self._p(moments), self._p(cdev), self._p(shift),
                                                                  self.p, C.byref(options), C.byref(out_struct),
                                                                  self._p(flag), self._p(rss), self._stream(),
                                                                  C.byref(err))
        self._check(rc, err)

    def residual(self, y, x, w, coefficients, intercept, shift, flag):
        """This shard's sum w (y - b0 - x'b)^2 (device f64[1]); zero, at no cost, when flag[0] == 0."""
        out = self.torch.zeros(1, dtype=self.torch.float64, device=self.device)
        err = abi.AnofoxError()
        rc = self.lib.anofox_b200_shard_residual_device(self._p(y), self._p(x), self._p(w), y.numel(), self.p,
                                                        self._p(coefficients), self._p(intercept), self._p(shift),
                                                        self._p(flag), self._p(out), C.c_void_p(self.ws_ptr), self.ws_bytes,
                                                        self._stream(), C.byref(err))
        self._check(rc, err)
        return out


def choose_shift(gathered):
    """gathered: [world, p + 2] rows (found, x_1..x_p, y). Returns the shift of the lowest rank with found = 1
    (zeros if no rank has a valid row: the reduced block then says n_valid = 0 -> NO_VALID_DATA)."""
    for r in range(gathered.shape[0]):
        if float(gathered[r, 0]) != 0.0:
            return gathered[r, 1:].clone()
    return gathered[0, 1:].clone() * 0.0


def sharded_moments(backend, y, x, w, dist=None, group=None):
    """Steps 0-2 on this rank's rows. Returns (moments, cdev, shift), identical on every rank."""
    cand = backend.first_valid(y, x, w)
    if dist is not None and dist.is_initialized() and dist.get_world_size(group) > 1:
        world = dist.get_world_size(group)
        bufs = [cand.new_empty(cand.shape) for _ in range(world)]
        dist.all_gather(bufs, cand, group=group)
        gathered = cand.new_empty((world,) + tuple(cand.shape))
        for r in range(world):
            gathered[r] = bufs[r]
    else:
        gathered = cand.reshape(1, -1)
    shift = choose_shift(gathered).contiguous()
    m, cd = backend.moments(y, x, w, shift)
    if dist is not None and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(m, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(cd, op=dist.ReduceOp.MAX, group=group)
    return m, cd, shift


class ShardedSingleGroupFit:
    """ols/ridge/wls fit of one group whose rows live on several GPUs (one shard per rank)."""

    def __init__(self, n_rows_local_max, p, inference=True, device="cuda", group=None):
        import torch
        import torch.distributed as dist
        from . import DeviceFit
        self.dist, self.group = dist, group
        self.backend = CudaBackend(n_rows_local_max, p, torch.device(device))
        self.fit_out = DeviceFit(1, 1, p, inference=inference, device=device)
        self.flag = torch.zeros(1, dtype=torch.int32, device=device)

    def fit(self, y, x, w=None, **opt):
        o = abi.make_fit_options(**opt)
        m, cd, shift = sharded_moments(self.backend, y, x, w, self.dist, self.group)
        out = self.fit_out.out
        self.backend.solve(m, cd, shift, o, self.fit_out.out_struct, flag=self.flag)
        # exact-fit guard (no host read of the flag: the kernels look at it themselves)
        rss = self.backend.residual(y, x, w, out["coefficients"], out["intercept"], shift, self.flag)
        dist = self.dist
        if dist is not None and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(rss, op=dist.ReduceOp.SUM, group=self.group)
        self.backend.solve(m, cd, shift, o, self.fit_out.out_struct, flag=self.flag, rss=rss)
        return out


class MixedShardedFit:
    """GROUP BY over several GPUs where a few groups are too large for one rank (SURVEY.md section 8e "mixed
    case"): sharding.mixed_plan decides, per rank, which whole groups it fits (no collective) and which row slice of
    every giant group it accumulates (one SUM + one MAX all-reduce per giant, as ShardedSingleGroupFit).

    fit() takes the CSR of the whole query (host offsets) and this rank's view of the rows as device tensors
    indexed by GLOBAL row (test scale: the full arrays; in production only plan.rows a rank touches need to be
    resident). Results land in full-size [G] arrays: the rank's own groups and every giant group are valid."""

    def __init__(self, offsets, world, rank, p, giant_rows, inference=True, device="cuda", group=None):
        import numpy as np
        import torch
        import torch.distributed as dist
        from . import DeviceFit
        from . import sharding
        self.torch, self.dist, self.group, self.rank, self.p = torch, dist, group, rank, p
        self.offsets = np.asarray(offsets, dtype=np.int64)
        self.plan = sharding.mixed_plan(self.offsets, world, giant_rows)
        G = len(self.offsets) - 1
        self.full = DeviceFit(G, 1, p, inference=inference, device=device)  # output arrays only
        self.inference, self.device = inference, device
        self._DeviceFit = DeviceFit

    def fit(self, y, x, w=None, **opt):
        torch, off = self.torch, self.offsets
        out = self.full.out
        out["status"].fill_(-1)  # -1 = not this rank's group
        off_dev = torch.from_numpy(off).to(self.device)
        for g0, g1 in self.plan["segments"][self.rank]:
            r0, r1 = int(off[g0]), int(off[g1])
            seg = self._DeviceFit(g1 - g0, r1 - r0, self.p, inference=self.inference, device=self.device)
            loc = (off_dev[g0:g1 + 1] - r0).contiguous()
            res = seg.fit(loc, y[r0:r1], x[r0:r1], None if w is None else w[r0:r1], **opt)
            for k, v in res.items():
                out[k][g0:g1] = v
        for g in self.plan["giants"]:
            a, b = self.plan["giant_rows"][g][self.rank]
            one = ShardedSingleGroupFit(max(b - a, 1), self.p, inference=self.inference, device=self.device,
                                        group=self.group)
            res = one.fit(y[a:b], x[a:b], None if w is None else w[a:b], **opt)
            for k, v in res.items():
                out[k][g:g + 1] = v
        return out
