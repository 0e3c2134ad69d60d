"""
Tensor-level wrappers around the C ABI of libpctts.so (include/pctts.h).

Every method takes torch CUDA tensors (used only as typed device memory), derives raw pointers / leading dimensions
from the views, and launches the hand-written sm_100a kernel on torch's current CUDA stream.  No method has a
PyTorch or CPU fallback: constructing `CudaOps` without a B200 and a built libpctts.so raises.

The engine (percivaltts_b200/engine.py) is written against this interface; tests/cpu_ops_double.py implements the
same interface with plain torch on the CPU *for testing the host logic only* (it is not importable from the package).
"""
import contextlib
import ctypes
import os
import weakref

import torch

from . import _lib

EPI_BIAS, EPI_LRELU, EPI_SIGMOID, EPI_DLRELU_AUX, EPI_ACCUM, EPI_ATOMIC, EPI_ADDMOD, EPI_WFLIP = 1, 2, 4, 8, 16, 32, 64, 128


def _mat(t):
    """(ptr, ld, seg, segstride, rows, cols) of a 2-D [R,C] or segmented 3-D [S,R,C] fp32 view, unit inner stride."""
    assert t.dtype == torch.float32, t.dtype
    if t.dim() == 2:
        R, C = t.shape
        assert C == 1 or t.stride(1) == 1, ('inner stride', t.stride())
        ld = t.stride(0) if R > 1 else max(C, 1)
        return t.data_ptr(), ld, 0, 0, R, C
    assert t.dim() == 3, t.shape
    S, R, C = t.shape
    assert C == 1 or t.stride(2) == 1, ('inner stride', t.stride())
    ld = t.stride(1) if R > 1 else max(C, 1)
    return t.data_ptr(), ld, R, t.stride(0), S * R, C


def _map4(t):
    """(ptr, batch stride, row pitch) of an NHWC map view [N,T,F,C] whose (F,C) plane is dense."""
    assert t.dtype == torch.float32 and t.dim() == 4, (t.dtype, t.shape)
    N, T, F, C = t.shape
    assert (C == 1 or t.stride(3) == 1) and (F == 1 or t.stride(2) == C), ('map strides', t.shape, t.stride())
    return t.data_ptr(), t.stride(0), t.stride(1)


def _ptr(t):
    return None if t is None else t.data_ptr()


class ReduceWorkspace(object):
    """Scratch of the deterministic two-stage column reductions; `sums` is the fp64 [2,C] result (all-reducible)."""

    def __init__(self, lib, rows, C, device):
        nbytes = lib.pctts_colreduce_workspace_bytes(rows, C)
        self.buf = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        off = lib.pctts_colreduce_sums_ptr(self.buf.data_ptr(), rows, C) - self.buf.data_ptr()
        self.sums = self.buf[off:off + 16 * C].view(torch.float64).view(2, C)
        self.rows, self.C = rows, C

    @property
    def ptr(self):
        return self.buf.data_ptr()


def _extent(t):
    """[lo, hi) byte range touched by the view t."""
    lo = t.data_ptr()
    n = 1
    for size, stride in zip(t.shape, t.stride()):
        if size == 0:
            return lo, lo
        n += (size - 1) * stride
    return lo, lo + n * t.element_size()


class Shadow(object):
    """bf16 [hi | lo] planes of an fp32 matrix view, the operand format of the tcgen05 GEMMs (pctts_split_bf16):
    `parts` planes of [rows, ld] bf16, ld = cols rounded up to 8 elements (16-byte TMA row pitch)."""

    def __init__(self, device, rows, cols, parts):
        self.rows, self.cols, self.parts = int(rows), int(cols), int(parts)
        self.ld = (self.cols + 7) // 8 * 8
        self.part_stride = self.rows * self.ld
        self.buf = torch.empty(self.parts * self.part_stride, dtype=torch.bfloat16, device=device)

    @property
    def ptr(self):
        return self.buf.data_ptr()


class ShadowRows(object):
    """Rows [r0, r0 + rows) of a parent Shadow, addressed in place (same pitch, same plane stride)."""

    def __init__(self, parent, r0, rows):
        self.parent, self.r0, self.rows = parent, int(r0), int(rows)
        self.cols, self.parts, self.ld, self.part_stride = parent.cols, parent.parts, parent.ld, parent.part_stride

    @property
    def ptr(self):
        return self.parent.ptr + 2 * self.r0 * self.ld


class _RowParent(object):
    """A contiguous [R, C] activation matrix whose row ranges are used as GEMM operands on their own (the batched critic
    applications): one shadow for the whole matrix, freshness tracked per row interval."""

    def __init__(self, t, shadow):
        self.t, self.shadow = t, shadow
        self.lo, self.hi = _extent(t)
        self.rowbytes = t.stride(0) * 4
        self.fresh = []            # disjoint sorted [a, b) row intervals that are up to date

    def missing(self, a, b):
        out, cur = [], a
        for x, y in self.fresh:
            if y <= cur:
                continue
            if x >= b:
                break
            if x > cur:
                out.append((cur, x))
            cur = max(cur, y)
        if cur < b:
            out.append((cur, b))
        return out

    def add(self, a, b):
        iv = sorted(self.fresh + [(a, b)])
        merged = [iv[0]]
        for x, y in iv[1:]:
            if x <= merged[-1][1]:
                merged[-1] = (merged[-1][0], max(merged[-1][1], y))
            else:
                merged.append((x, y))
        self.fresh = merged

    def drop(self, a, b):
        out = []
        for x, y in self.fresh:
            if y <= a or x >= b:
                out.append((x, y))
            else:
                if x < a:
                    out.append((x, a))
                if y > b:
                    out.append((b, y))
        self.fresh = out


class ChainLayerStruct(ctypes.Structure):
    """pctts_chain_layer of include/pctts.h"""
    _fields_ = [('w', ctypes.c_void_p), ('bias', ctypes.c_void_p), ('aux', ctypes.c_void_p), ('y', ctypes.c_void_p),
                ('ybatch', ctypes.c_long), ('ypitch', ctypes.c_long), ('auxbatch', ctypes.c_long), ('auxpitch', ctypes.c_long),
                ('cin', ctypes.c_int), ('cout', ctypes.c_int), ('flags', ctypes.c_int), ('flip', ctypes.c_int)]


class _AblatedLib(object):
    """libpctts.so with some entry points stubbed out (PCTTS_ABLATE=pctts_conv2d_chain,...): how much of a step's time a kernel
    family is responsible for, measured by leaving it out.  Never on in a run whose results are used."""

    def __init__(self, lib, names):
        self._lib, self._names = lib, set(n.strip() for n in names if n.strip())

    def __getattr__(self, name):
        if name in self._names:
            return lambda *a, **k: 0
        return getattr(self._lib, name)


class SplitJobStruct(ctypes.Structure):
    """pctts_split_job of include/pctts.h"""
    _fields_ = [('src', ctypes.c_void_p), ('dst', ctypes.c_void_p), ('ld', ctypes.c_long), ('rows', ctypes.c_long),
                ('ld_dst', ctypes.c_long), ('part_stride', ctypes.c_long), ('cols', ctypes.c_int)]


class WgradLayerStruct(ctypes.Structure):
    """pctts_wgrad_layer of include/pctts.h"""
    _fields_ = [('x', ctypes.c_void_p), ('dy', ctypes.c_void_p), ('dw', ctypes.c_void_p), ('dbias', ctypes.c_void_p),
                ('xbatch', ctypes.c_long), ('xpitch', ctypes.c_long), ('dybatch', ctypes.c_long), ('dypitch', ctypes.c_long),
                ('cin', ctypes.c_int)]


PRECISIONS = {'fp32': 6, 'fp32x3': 3, 'bf16': 1, 'simt': 0}      # -> tcgen05.mma products per k-slice
PLANES = {6: 3, 3: 2, 1: 1, 0: 0}


class CudaOps(object):
    name = 'cuda'

    def __init__(self, device=None, precision=None):
        """precision of the dense contractions (everything else is fp32 CUDA-core arithmetic):
        'fp32'   (default) tcgen05 GEMMs on three-plane bf16 expansions (24 significand bits), the six products of order
                 <= 2 per k-slice, fp32 accumulation in TMEM: fp32-grade results (products exact to ~2^-23);
        'fp32x3' two planes / three products: products exact to ~2^-16 -- faster, inside 1e-3 on outputs and losses but
                 not on every gradient of an ill-conditioned (tiny batch-norm batch) step;
        'bf16'   one plane / one product (BASELINE config #3; tolerance stated separately);
        'simt'   exact fp32 CUDA-core GEMMs (no tensor cores)."""
        if not torch.cuda.is_available():
            raise _lib.PcttsError('percivaltts_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback')
        self.lib = _lib.load()
        if os.environ.get('PCTTS_ABLATE'):      # developer probe: named C-ABI calls become no-ops (timing ablations; results are wrong)
            self.lib = _AblatedLib(self.lib, os.environ['PCTTS_ABLATE'].split(','))
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        _lib.check(self.lib.pctts_device_check(self.device.index or 0), 'pctts_device_check')
        precision = precision or os.environ.get('PCTTS_PRECISION', 'fp32')
        if precision not in PRECISIONS:
            raise ValueError('unknown precision %r (one of %s)' % (precision, sorted(PRECISIONS)))
        self.precision = precision
        self.nterms = PRECISIONS[precision]
        # products per k-slice inside backward_terms(): the gradient passes of the critic are LINEAR in their seeds (LeakyReLU masks come
        # from the stored fp32 activations), so two planes / three products (2^-17 per product) leave the gradients ~1e-5 from the fp32
        # result; the forward passes, which decide the masks, keep the six products.  PCTTS_BWD_TERMS=6 switches it off.
        self.bwd_nterms = 3 if (self.nterms == 6 and os.environ.get('PCTTS_BWD_TERMS', '3') == '3') else self.nterms
        self._nt = self.nterms
        self.fuse_bn_stats = os.environ.get('PCTTS_FUSE_BN_STATS', '1') != '0'   # BN sums in the Conv2D epilogue
        self.bn_single_launch = os.environ.get('PCTTS_BN_SINGLE_LAUNCH', '1') != '0'   # second stage inside bn_finalize (1 replica)
        self.conv2d_tc = os.environ.get('PCTTS_CONV2D_TC', '1') != '0'   # tcgen05 implicit-GEMM Conv2D (4->4, 5x5)
        self.split_multi = os.environ.get('PCTTS_SPLIT_MULTI', '1') != '0'   # weight shadows after Adam in one launch
        self.fuse_planes = os.environ.get('PCTTS_FUSE_PLANES', '1') != '0'   # BatchNorm apply kernels write the bf16 operand planes of their output
        self.tc_wflip = os.environ.get('PCTTS_TC_WFLIP', '1') != '0'     # Conv2D data gradients read the forward kernels (no flip launch)
        # PCTTS_SPLIT_FWD=1: batched critic forward as two chain launches (18 + 12 samples), the post layers of the first under the second.
        # Measured on B200: critic step 1.930 -> 2.007 ms (GEMM CTAs on the free SMs delay the second launch's 8-CTA clusters) -- off
        self.split_fwd = os.environ.get('PCTTS_SPLIT_FWD', '0') != '0'
        self.wgrad_multi = os.environ.get('PCTTS_WGRAD_MULTI', '1') != '0'   # weight gradients of a whole Conv2D stack in one launch
        self.conv_chain = os.environ.get('PCTTS_CONV_CHAIN', '1') != '0'   # the critic's BN-free Conv2D stack as ONE launch per chain
        self._shadows = {}         # view key -> Shadow
        self._fresh = {}           # view key -> (lo, hi): activation shadows valid inside the current step scope
        self._persist_ranges = []  # (weakref of the parameter buffer, lo, hi): shadows of views into it are refreshed by
                                   # refresh_persistent(); a range dies with its buffer (the addresses may be reused)
        self._persist = {}         # view key -> (Shadow, view)
        self._tracking = False
        self._parents = []         # _RowParent list (share_rows)
        self._red = {}
        self._wg = {}
        self._lse_ws = None
        self._lse_ws_lane = {}     # lane -> workspace of the column-window LSE (calls on different lanes may overlap)
        self._splitk_ws = {}       # lane -> split-K workspace
        self._ws_keep = []
        self._tag = None           # batch-shape tag (shape_scope): per-shape scratch is recorded under it and freed by release_shape()
        self._tagged = {}          # tag -> list of (dict, key) entries created while the tag was active
        self._lane = 0             # 0 = the caller's stream; 1 = the side stream of side_lane()
        self._side = None
        self._side2 = None         # second side stream (side_lane(second=True)), joined by join_lanes()
        self._side2_open = False
        self.lane3 = os.environ.get('PCTTS_LANE3', '1') != '0'
        self.lanes = os.environ.get('PCTTS_LANES', '1') != '0'   # run independent branches of a step on two streams
        self.timeline = None
        self.launches = 0          # number of C-ABI kernel-launching calls (reset by the caller)

    # -- helpers ---------------------------------------------------------------------------------------------
    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _ck(self, rc, what):
        self.launches += 1
        _lib.check(rc, what)
        if self.timeline is not None:      # developer probe (tools/timeline.py): a time stamp behind every call, per lane
            i = len(self.timeline)
            if i < self._stamps.numel():
                self.lib.pctts_stamp(self._stamps.data_ptr() + 8 * i, self._stream())
                self.timeline.append((what, self._lane))

    def timeline_start(self, n=4096):
        self._stamps = torch.zeros(n, dtype=torch.int64, device=self.device)
        self.timeline = []
        self.stamp('start')

    def stamp(self, what):
        if self.timeline is not None:
            self.lib.pctts_stamp(self._stamps.data_ptr() + 8 * len(self.timeline), self._stream())
            self.timeline.append((what, self._lane))

    def timeline_stop(self):
        """-> [(what, lane, ns since start)]"""
        torch.cuda.synchronize(self.device)
        t = self._stamps.cpu().tolist()
        out = [(w, l, t[i] - t[0]) for i, (w, l) in enumerate(self.timeline)]
        self.timeline = None
        return out

    def reduce_ws(self, rows, C):
        key = (self._tag, self._lane, int(rows), int(C))
        ws = self._red.get(key)
        if ws is None:
            ws = self._red[key] = ReduceWorkspace(self.lib, key[2], key[3], self.device)
            self._note(self._red, key)
        return ws

    # -- per-batch-shape scratch (operand shadows of activations, reduction / weight-gradient workspaces, row parents) --------
    @contextlib.contextmanager
    def shape_scope(self, tag):
        """Scratch allocated inside belongs to the batch shape `tag` and is freed as a whole by release_shape(tag): the step's
        buffers are created per (B, T), and a training run meets many T (shortest utterance of each batch)."""
        outer, self._tag = self._tag, tag
        try:
            yield self
        finally:
            self._tag = outer

    def _note(self, table, key):
        if self._tag is not None:
            self._tagged.setdefault(self._tag, []).append((table, key))

    def release_shape(self, tag):
        """Free everything recorded under `tag` (captured graphs that used it must be dropped by the caller first)."""
        for table, key in self._tagged.pop(tag, []):
            if isinstance(table, list):
                if key in table:
                    table.remove(key)
            else:
                table.pop(key, None)
        self._fresh.clear()

    # -- two-lane execution: independent branches of a step on two streams (fork / join, capturable) ------------------
    @contextlib.contextmanager
    def side_lane(self, second=False, lane=None, nowait=False):
        """Kernels issued inside run on a side stream, ordered after everything issued so far on the caller's stream, and
        concurrently with what the caller issues afterwards until join_lanes().  Scratch workspaces are per lane; operands
        shared by lanes must have their bf16 shadows taken BEFORE the fork (shadow()).  Two side streams exist: lane 1 (default,
        forked from the caller's stream) and lane 2 (lane=2 from anywhere, or second=True inside lane 1: a branch off the first
        side lane; PCTTS_LANE3=0 runs it in line).  With PCTTS_LANES=0 the body simply runs in line."""
        if lane is None:
            lane = 2 if (second and self._lane == 1) else 1
        if not self.lanes or lane == self._lane or (lane == 1 and self._lane != 0) or (lane == 2 and not self.lane3):
            yield self
            return
        cur = torch.cuda.current_stream(self.device)
        if lane == 2:
            if self._side2 is None:
                self._side2 = torch.cuda.Stream(device=self.device)
            side = self._side2
            self._side2_open = True
        else:
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.device)
            side = self._side
        if not nowait:      # nowait: continue an already forked lane where it stands (what is queued needs nothing newer of the caller's lane)
            side.wait_stream(cur)
        outer, self._lane = self._lane, lane
        try:
            with torch.cuda.stream(side):
                yield self
        finally:
            self._lane = outer

    @contextlib.contextmanager
    def backward_terms(self):
        """Tensor-core contractions issued inside use the gradient-pass precision (see bwd_nterms); operand planes are unchanged
        (the two-plane products read the first two of the three planes)."""
        outer, self._nt = self._nt, self.bwd_nterms
        try:
            yield self
        finally:
            self._nt = outer

    @contextlib.contextmanager
    def full_terms(self):
        """Inside backward_terms(): back to the forward precision for a pass whose result is differenced afterwards (the input
        gradient of the interpolated rows feeds 1 - ||g||)."""
        outer, self._nt = self._nt, self.nterms
        try:
            yield self
        finally:
            self._nt = outer

    def preshare(self, t):
        """Take the bf16 shadow of an operand both lanes are going to read (no-op for the CUDA-core GEMM mode)."""
        if self.nterms > 0:
            self.shadow(t)

    def lane_event(self):
        """Inside a side lane: marks the point reached there; another lane can wait for it without a full join."""
        if not self.lanes or self._lane == 0:
            return None
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        return ev

    def wait_lane_event(self, ev):
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)

    def join_lanes(self, which=None):
        """The caller's stream waits for the side lanes (which = 1 / 2: for that one only)."""
        if self.lanes and self._side is not None and which != 2:
            torch.cuda.current_stream(self.device).wait_stream(self._side)
        if self._side2_open and which != 1:
            torch.cuda.current_stream(self.device).wait_stream(self._side2)
            self._side2_open = False

    def empty(self, *shape, dtype=torch.float32):
        return torch.empty(*shape, dtype=dtype, device=self.device)

    def zeros(self, *shape, dtype=torch.float32):
        return torch.zeros(*shape, dtype=dtype, device=self.device)

    # -- bf16 operand shadows of the tensor-core GEMMs ----------------------------------------------------------
    @contextlib.contextmanager
    def step_scope(self):
        """Inside the scope an activation is split once and its shadow reused until a kernel of this object overwrites the
        memory it came from (every writer reports through _wrote).  Outside a scope every GEMM re-splits its operands."""
        outer = self._tracking
        if not outer:
            self._clear_fresh()
        self._tracking = True
        try:
            yield self
        finally:
            self._tracking = outer
            if not outer:
                self._clear_fresh()

    def register_persistent(self, flat):
        """Declare a parameter buffer: shadows of views into it stay valid until refresh_persistent(flat)."""
        lo, hi = _extent(flat)
        self._persist_ranges.append((weakref.ref(flat), lo, hi))

    def _is_persistent(self, ptr):
        """ptr lies in a LIVE registered parameter buffer.  Ranges of buffers that no longer exist are dropped together with the
        shadows taken from them: their memory goes back to the allocator and may come back as an activation."""
        dead = [r for r in self._persist_ranges if r[0]() is None]
        if dead:
            self._persist_ranges = [r for r in self._persist_ranges if r[0]() is not None]
            for _, lo, hi in dead:
                for key in [k for k in self._persist if lo <= k[0] < hi]:
                    del self._persist[key]
                    self._shadows.pop(key, None)
        return any(lo <= ptr < hi for _, lo, hi in self._persist_ranges)

    def refresh_persistent(self, flat):
        """Re-split every shadow taken from views of `flat` (after Adam or a weight load)."""
        lo, hi = _extent(flat)
        jobs = []
        for key, (sh, view) in self._persist.items():
            if lo <= key[0] < hi:
                p, ld, seg, ss, rows, cols = _mat(view)
                if seg > 0 or not self.split_multi:
                    self._split(view, sh)
                else:
                    jobs.append(SplitJobStruct(p, sh.ptr, ld, rows, sh.ld, sh.part_stride, cols))
        for i in range(0, len(jobs), 24):       # the plain matrices in one launch per 24
            chunk = jobs[i:i + 24]
            arr = (SplitJobStruct * len(chunk))(*chunk)
            self._ck(self.lib.pctts_split_bf16_multi(ctypes.cast(arr, ctypes.c_void_p), len(chunk), PLANES[self.nterms], self._stream()),
                     'pctts_split_bf16_multi')

    def _clear_fresh(self):
        self._fresh.clear()
        for par in self._parents:
            par.fresh = []

    def share_rows(self, t):
        """Declare a contiguous [R, C] activation matrix whose row ranges are GEMM operands of their own: every such view
        then addresses rows of ONE shadow, and a row is split once per step no matter through which view it is used."""
        if self.nterms == 0:
            return
        assert t.dim() == 2 and t.is_contiguous()
        par = _RowParent(t, Shadow(self.device, t.shape[0], t.shape[1], PLANES[self.nterms]))
        self._parents.append(par)
        self._note(self._parents, par)

    def _row_parent(self, t):
        if t.dim() != 2 or not self._parents:
            return None, 0
        ptr = t.data_ptr()
        for par in self._parents:
            if par.lo <= ptr < par.hi and t.shape[1] == par.t.shape[1] and (t.shape[0] == 1 or t.stride(0) == par.t.stride(0)) \
                    and (ptr - par.lo) % par.rowbytes == 0:
                return par, (ptr - par.lo) // par.rowbytes
        return None, 0

    def _wrote(self, t):
        if t is None:
            return
        if self._fresh:
            lo, hi = _extent(t)
            for key in [k for k, (a, b) in self._fresh.items() if a < hi and lo < b]:
                del self._fresh[key]
        for par in self._parents:
            if par.fresh:
                lo, hi = _extent(t)
                if lo < par.hi and par.lo < hi:
                    a = max(lo - par.lo, 0) // par.rowbytes
                    b = (min(hi, par.hi) - par.lo + par.rowbytes - 1) // par.rowbytes
                    par.drop(a, b)

    def _split(self, t, sh):
        p, ld, seg, ss, rows, cols = _mat(t)
        self._ck(self.lib.pctts_split_bf16(p, ld, seg, ss, rows, cols, sh.ptr, sh.ld, sh.part_stride, sh.parts, self._stream()),
                 'pctts_split_bf16')

    def shadow(self, t):
        """Up-to-date bf16 shadow of the 2-D / segmented 3-D fp32 view t."""
        par, r0 = self._row_parent(t)
        if par is not None:
            r1 = r0 + t.shape[0]
            for a, b in (par.missing(r0, r1) if self._tracking else [(r0, r1)]):
                self._split(par.t[a:b], ShadowRows(par.shadow, a, b - a))
            if self._tracking:
                par.add(r0, r1)
            return ShadowRows(par.shadow, r0, t.shape[0])
        key = (t.data_ptr(), tuple(t.shape), tuple(t.stride()))
        sh = self._shadows.get(key)
        ptr = key[0]
        if sh is None:
            _, _, _, _, rows, cols = _mat(t)
            sh = self._shadows[key] = Shadow(self.device, rows, cols, PLANES[self.nterms])
            if not self._is_persistent(ptr):
                self._note(self._shadows, key)
        if self._is_persistent(ptr):
            if key not in self._persist:
                self._persist[key] = (sh, t)
                self._split(t, sh)
            return sh
        if key not in self._fresh:
            self._split(t, sh)
            if self._tracking:
                self._fresh[key] = _extent(t)
        return sh

    def _tc_ok(self, M, N, K, C, flags):
        return (self.nterms > 0 and C.dim() == 2 and M >= 32 and N >= 8 and K >= 16 and not (flags & EPI_ATOMIC))

    # -- dense contractions ----------------------------------------------------------------------------------
    def gemm(self, A, B, C, transA=False, transB=False, alpha=1.0, bias=None, aux=None, addmod=None, flags=0):
        """C = epilogue(alpha * op(A) op(B)); A/B/C are 2-D or segmented 3-D views (see _mat)."""
        pa, lda, sa, ssa, ra, ca = _mat(A)
        pb, ldb, sb, ssb, rb, cb = _mat(B)
        pc, ldc, sc, ssc, rc_, cc = _mat(C)
        M, K = (ca, ra) if transA else (ra, ca)
        Kb, N = (cb, rb) if transB else (rb, cb)
        assert K == Kb and rc_ == M and cc == N, ('gemm shapes', A.shape, B.shape, C.shape, transA, transB)
        addrows = 0
        if bias is not None:
            flags |= EPI_BIAS
            assert bias.numel() == N and bias.is_contiguous()
        if aux is not None:
            flags |= EPI_DLRELU_AUX
            pa2, ldx, sx, ssx, rx, cx = _mat(aux)
            assert (ldx, sx, ssx, rx, cx) == (ldc, sc, ssc, rc_, cc), 'aux must share the addressing of C'
        if addmod is not None:
            flags |= EPI_ADDMOD
            assert addmod.dim() == 2 and addmod.shape[1] == N and addmod.is_contiguous()
            addrows = addmod.shape[0]
        if self._tc_ok(M, N, K, C, flags):
            sha, shb = self.shadow(A), self.shadow(B)
            self._wrote(C)
            splits, ws = 1, None
            if (flags & ~EPI_ACCUM) == 0:
                splits = self.lib.pctts_gemm_tc_splits(M, N, K)
                if splits > 1:
                    ws = self._splitk(self.lib.pctts_gemm_tc_workspace(M, N, splits))
            if splits == 1 and (flags & ~(EPI_BIAS | EPI_LRELU | EPI_DLRELU_AUX | EPI_ACCUM)) == 0:
                need = self.lib.pctts_gemm_tc_tail_workspace(M, N, K)     # more tiles than SMs: K-split of the last wave's tiles
                if need:
                    ws = self._splitk(need)
                    self.launches += 1
            self._ck(self.lib.pctts_gemm_tc(int(transA), int(transB), M, N, K, float(alpha), sha.ptr, sha.ld, sha.part_stride,
                                            shb.ptr, shb.ld, shb.part_stride, pc, ldc, _ptr(bias), _ptr(aux), _ptr(addmod),
                                            addrows, flags, self._nt, splits, _ptr(ws), self._stream()), 'pctts_gemm_tc')
            return
        self._wrote(C)
        if (flags & ~EPI_ACCUM) == 0 and K >= 1024:
            # long reduction, few output tiles (weight gradients): split K over the idle SMs, deterministic two-stage sum
            tiles = ((M + 127) // 128) * ((N + 63) // 64)
            splits = min((2 * 148) // tiles, K // 256, 64)
            if splits > 1:
                ws = self._splitk(splits * M * N)
                self._ck(self.lib.pctts_gemm_splitk_f32(int(transA), int(transB), M, N, K, float(alpha), pa, lda, sa, ssa, pb, ldb,
                                                        sb, ssb, pc, ldc, sc, ssc, flags, splits, ws.data_ptr(),
                                                        self._stream()), 'pctts_gemm_splitk_f32')
                return
        self._ck(self.lib.pctts_gemm_f32(int(transA), int(transB), M, N, K, float(alpha), pa, lda, sa, ssa, pb, ldb, sb, ssb,
                                         pc, ldc, sc, ssc, _ptr(bias), _ptr(aux), _ptr(addmod), addrows, flags,
                                         self._stream()), 'pctts_gemm_f32')

    def _splitk(self, need):
        ws = self._splitk_ws.get(self._lane)
        if ws is None or ws.numel() < need:
            self._ws_keep.append(ws)                   # captured graphs may still point at the old one
            ws = self._splitk_ws[self._lane] = self.empty(max(int(need), 1))
        return ws

    # -- 1-D convolutions (pCNN1D, networktts.py:116-120) --------------------------------------------------------
    def conv1d(self, xp, kernel, y, T, bias=None, aux=None, flags=0):
        """'same' Conv1D of the time-padded input xp [B, T+W-1, CIN] with the keras kernel [W, CIN, COUT] -> y [B*T, COUT]
        (virtual im2col: no patch matrix is ever written)."""
        B, Tp, CIN = xp.shape
        W, ci, COUT = kernel.shape
        assert ci == CIN and Tp == T + W - 1 and xp.is_contiguous() and kernel.is_contiguous()
        assert tuple(y.shape) == (B * T, COUT) and y.stride(1) == 1
        if self.nterms > 0 and COUT >= 8:
            if bias is not None:
                flags |= EPI_BIAS
            if aux is not None:
                flags |= EPI_DLRELU_AUX
                assert tuple(aux.shape) == tuple(y.shape) and aux.stride() == y.stride()
            shx, shw = self.shadow(xp.view(B * Tp, CIN)), self.shadow(kernel.view(W * CIN, COUT))
            self._wrote(y)
            need = self.lib.pctts_conv1d_tc_fwd_workspace(B, T, W, CIN, COUT)      # two K halves per tile when the tiling is short of CTAs
            ws = self._splitk(need) if need else None
            if ws is not None:
                self.launches += 1
            self._ck(self.lib.pctts_conv1d_tc_fwd(shx.ptr, shx.ld, shx.part_stride, shw.ptr, shw.ld, shw.part_stride, y.data_ptr(),
                                                  y.stride(0), B, T, W, CIN, COUT, _ptr(bias), _ptr(aux), flags, self._nt,
                                                  _ptr(ws), self._stream()), 'pctts_conv1d_tc_fwd')
            return
        self.gemm(xp.as_strided((B, T, W * CIN), (Tp * CIN, CIN, 1)), kernel.view(W * CIN, COUT), y, bias=bias, aux=aux, flags=flags)

    def conv1d_wgrad(self, xp, dz, dw, T):
        """dw [W, CIN, COUT] = sum_{b,t} xp[b, t+w, :]^T dz[b*T+t, :]"""
        B, Tp, CIN = xp.shape
        W, ci, COUT = dw.shape
        assert ci == CIN and Tp == T + W - 1 and xp.is_contiguous() and dw.is_contiguous()
        assert tuple(dz.shape) == (B * T, COUT) and dz.is_contiguous()
        if self.nterms > 0 and COUT >= 8:
            shx, shz = self.shadow(xp.view(B * Tp, CIN)), self.shadow(dz)
            need = self.lib.pctts_conv1d_tc_wgrad_workspace(B, T, W, CIN, COUT)
            ws = self._splitk(need) if need else None
            if ws is not None:
                self.launches += 1
            self._ck(self.lib.pctts_conv1d_tc_wgrad(shx.ptr, shx.ld, shx.part_stride, shz.ptr, shz.ld, shz.part_stride,
                                                    dw.data_ptr(), B, T, W, CIN, COUT, 0, self._nt, _ptr(ws), self._stream()),
                     'pctts_conv1d_tc_wgrad')
            return
        self.gemm(xp.as_strided((B, T, W * CIN), (Tp * CIN, CIN, 1)), dz, dw.view(W * CIN, COUT), transA=True)

    # -- 2-D convolutions ------------------------------------------------------------------------------------
    def conv2d_dgrad(self, dy, w, dx, scratch, aux=None, flags=0):
        """dx = epilogue(conv2d_same(dy, flip(w))): the data gradient of y = conv2d_same(x, w), w the FORWARD kernel
        [KH,KW,Cin,Cout].  The tensor-core kernel forms the flipped kernel on the fly; otherwise it goes to `scratch` [KH,KW,Cout,Cin]."""
        N, T, F, COUT = dy.shape
        KH, KW, CIN, co = w.shape
        assert co == COUT and tuple(dx.shape) == (N, T, F, CIN) and w.is_contiguous()
        px, xb, xp = _map4(dy)
        py, yb, yp = _map4(dx)
        pa, ab, ap = _map4(aux) if aux is not None else (0, 0, 0)
        if (self.nterms > 0 and self.conv2d_tc and self.tc_wflip and self.lib.pctts_conv2d_tc_supported(F, COUT, CIN, KH, KW, (KH - 1) // 2, (KW - 1) // 2)
                and not ((px | py | pa) & 15) and not ((xp | yp | ap | xb | yb | ab) & 3)):
            if aux is not None:
                flags |= EPI_DLRELU_AUX
            self._wrote(dx)
            self._ck(self.lib.pctts_conv2d_tc_fwd_fused(px, w.data_ptr(), None, pa or None, py, N, T, F, COUT, CIN, KH, KW, (KH - 1) // 2,
                                                        (KW - 1) // 2, xp, yp, ap, xb, yb, ab, flags | EPI_WFLIP, self._nt, None, None,
                                                        None, None, None, 1, self._stream()), 'pctts_conv2d_tc_fwd_fused')
            return
        self.conv2d_flip(w, scratch)
        self.conv2d(dy, scratch, dx, aux=aux, flags=flags)

    def conv2d(self, x, w, y, bias=None, aux=None, pad=None, flags=0, stats=False, in_bn=None, stats_final=True):
        """y = epilogue(conv2d_same(x, w)); x [N,T,F,Cin], w [KH,KW,Cin,Cout] contiguous, y/aux [N,T,F,Cout] views.
        stats=True also returns the ReduceWorkspace holding the BatchNorm sums [sum y, sum y^2] of the (contiguous) output:
        fused into the tensor-core kernel's epilogue where that kernel applies, a bn_stats() pass otherwise.
        in_bn = (mean, invstd, gamma, beta, scratch): x is a pre-BatchNorm map and the convolution reads LeakyReLU(BN(x)) --
        formed on load by the tensor-core kernel (nothing is written), else materialised into `scratch` first.
        stats_final=False (single replica): the workspace is left with its first-stage partials (ws.pending), which
        bn_finalize() folds into its own launch."""
        N, T, F, CIN = x.shape
        KH, KW, ci, COUT = w.shape
        assert ci == CIN and tuple(y.shape) == (N, T, F, COUT) and w.is_contiguous()
        pad_t, pad_f = ((KH - 1) // 2, (KW - 1) // 2) if pad is None else pad
        px, xb, xp = _map4(x)
        py, yb, yp = _map4(y)
        pa, ab, ap = (None, 0, 0)
        if bias is not None:
            flags |= EPI_BIAS
        if aux is not None:
            flags |= EPI_DLRELU_AUX
            assert tuple(aux.shape) == tuple(y.shape)
            pa, ab, ap = _map4(aux)
        self._wrote(y)
        tc = (self.nterms > 0 and self.conv2d_tc and self.lib.pctts_conv2d_tc_supported(F, CIN, COUT, KH, KW, pad_t, pad_f)
              and not ((px | py | (pa or 0)) & 15) and not ((xp | yp | ap | xb | yb | ab) & 3))
        bnp = (None, None, None, None)
        if in_bn is not None:
            mean, invstd, gamma, beta, scratch = in_bn
            if tc and self.fuse_bn_stats:
                bnp = (mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(), beta.data_ptr())
            else:
                self.bn_apply(x.view(N * T * F, CIN), scratch.view(N * T * F, CIN), mean, invstd, gamma, beta, 0.0, 0, 1)
                x = scratch
                px, xb, xp = _map4(x)
        if tc:
            # implicit GEMM on the tensor cores (4 -> 4 channels, 5x5): forward, data gradient (flipped kernels), GP tangent sweep
            if stats and self.fuse_bn_stats and y.is_contiguous():
                ws = self.reduce_ws(N * T * F, COUT)
                rc = self.lib.pctts_conv2d_tc_fwd_fused(px, w.data_ptr(), _ptr(bias), pa, py, N, T, F, CIN, COUT, KH, KW, pad_t,
                                                        pad_f, xp, yp, ap, xb, yb, ab, flags, self._nt, bnp[0], bnp[1], bnp[2],
                                                        bnp[3], ws.ptr, int(stats_final), self._stream())
                if rc != -2:      # PCTTS_EUNSUPPORTED: map too small for 148 partial slots
                    self.launches += int(stats_final)      # conv + the second reduction stage
                    self._ck(rc, 'pctts_conv2d_tc_fwd_fused')
                    ws.pending = 0 if stats_final else self.lib.pctts_conv2d_tc_stat_partials(N, T, F)
                    return ws
            self._ck(self.lib.pctts_conv2d_tc_fwd_fused(px, w.data_ptr(), _ptr(bias), pa, py, N, T, F, CIN, COUT, KH, KW, pad_t, pad_f,
                                                        xp, yp, ap, xb, yb, ab, flags, self._nt, bnp[0], bnp[1], bnp[2], bnp[3],
                                                        None, 1, self._stream()), 'pctts_conv2d_tc_fwd_fused')
        else:
            self._ck(self.lib.pctts_conv2d_fwd(px, w.data_ptr(), _ptr(bias), pa, py, N, T, F, CIN, COUT, KH, KW, pad_t, pad_f,
                                               xp, yp, ap, xb, yb, ab, flags, self._stream()), 'pctts_conv2d_fwd')
        if stats:
            return self.bn_stats(y.view(N * T * F, COUT), final=stats_final)
        return None

    def conv2d_chain_ok(self, T, F, KH, KW, channels):
        """The multi-layer chain kernel (conv2d_chain.cu) applies: 5x5 kernels, 4 feature channels, a precision with tensor-core
        operands, and a sequence length that fits 16 CTAs of shared memory per sample."""
        return bool(self.conv_chain and self.conv2d_tc and self.nterms > 0 and KH == 5 and KW == 5 and channels == 4
                    and self.lib.pctts_conv2d_chain_supported(int(T), int(F), self.nterms) > 0)

    def conv2d_chain(self, x, layers):
        """A chain of 'same' 5x5 Conv2D layers in one launch.  x [N,T,F,cin0]; layers: list of dicts with
        w (keras kernel [5,5,cin,cout], or the FORWARD kernel [5,5,cout,cin] with flip=True: the data-gradient kernel is formed
        on the fly), y [N,T,F,cout], optional bias, optional aux (LeakyReLU' mask source [N,T,F,4], may be y itself), flags."""
        N, T, F, cin0 = x.shape
        assert x.dtype == torch.float32 and (cin0 == 1 or x.stride(3) == 1) and (F == 1 or x.stride(2) == cin0)
        arr = (ChainLayerStruct * len(layers))()
        cin = cin0
        for i, L in enumerate(layers):
            w, y, flip = L['w'], L['y'], bool(L.get('flip', False))
            KH, KW, d2, d3 = w.shape
            cout = d2 if flip else d3
            assert KH == 5 and KW == 5 and w.is_contiguous() and (d3 if flip else d2) == cin, ('chain kernel', w.shape, cin, flip)
            assert tuple(y.shape) == (N, T, F, cout), (y.shape, (N, T, F, cout))
            py, yb, yp = _map4(y)
            flags = int(L.get('flags', 0))
            bias, aux = L.get('bias'), L.get('aux')
            pa, ab, ap = (None, 0, 0)
            if bias is not None:
                flags |= EPI_BIAS
                assert bias.numel() == cout and bias.is_contiguous()
            if aux is not None:
                flags |= EPI_DLRELU_AUX
                assert tuple(aux.shape) == (N, T, F, 4)
                pa, ab, ap = _map4(aux)
            self._wrote(y)
            arr[i] = ChainLayerStruct(w.data_ptr(), _ptr(bias), pa, py, yb, yp, ab, ap, cin, cout, flags, int(flip))
            cin = cout
        self._ck(self.lib.pctts_conv2d_chain(x.data_ptr(), x.stride(0), x.stride(1), cin0, N, T, F, len(layers),
                                             ctypes.cast(arr, ctypes.c_void_p), self._nt, self._stream()), 'pctts_conv2d_chain')

    def conv2d_flip(self, w, wt):
        KH, KW, CIN, COUT = w.shape
        assert tuple(wt.shape) == (KH, KW, COUT, CIN) and w.is_contiguous() and wt.is_contiguous()
        self._ck(self.lib.pctts_conv2d_flip_weights(w.data_ptr(), wt.data_ptr(), KH, KW, CIN, COUT, self._stream()),
                 'pctts_conv2d_flip_weights')

    def conv2d_wgrad(self, x, dy, dw, scale=1.0, accumulate=False, pad=None, dbias=None, bias_samples=None):
        """dw = scale * sum x (*) dy; with `dbias` also the bias gradient sum_{n < bias_samples} dy (same pass for the
        4 -> 4 channel 5x5 layers, a separate column sum otherwise)."""
        N, T, F, CIN = x.shape
        KH, KW, ci, COUT = dw.shape
        assert ci == CIN and tuple(dy.shape) == (N, T, F, COUT) and dw.is_contiguous()
        pad_t, pad_f = ((KH - 1) // 2, (KW - 1) // 2) if pad is None else pad
        key = (N, T, CIN, COUT, KH, KW)
        wkey = (self._tag, self._lane) + key
        ws = self._wg.get(wkey)
        if ws is None:
            ws = self._wg[wkey] = self.empty(max(self.lib.pctts_conv2d_wgrad_workspace(*key), 148 * (KH * KW * CIN * COUT + COUT)))
            self._note(self._wg, wkey)
        px, xb, xp = _map4(x)
        pd, db, dp = _map4(dy)
        nb = N if bias_samples is None else int(bias_samples)
        if dbias is not None:
            assert dbias.numel() == COUT and dbias.is_contiguous()
            rc = self.lib.pctts_conv2d_wgrad_bias(px, pd, dw.data_ptr(), dbias.data_ptr(), nb, ws.data_ptr(), N, T, F, CIN, COUT, KH, KW,
                                                  pad_t, pad_f, xp, dp, xb, db, float(scale), int(accumulate), self._stream())
            if rc != -2:
                self._ck(rc, 'pctts_conv2d_wgrad_bias')
                return
            assert dy.is_contiguous()
            self.colsum(dy.view(N * T * F, COUT)[:nb * T * F], dbias)
        self._ck(self.lib.pctts_conv2d_wgrad(px, pd, dw.data_ptr(), ws.data_ptr(), N, T, F, CIN, COUT, KH, KW, pad_t, pad_f,
                                             xp, dp, xb, db, float(scale), int(accumulate), self._stream()),
                 'pctts_conv2d_wgrad')

    def conv2d_wgrad_multi_ok(self, layers):
        """The one-launch weight gradient of a whole 5x5 stack applies (4 output channels, 1 or 4 input channels, aligned maps)."""
        if not self.wgrad_multi or not layers or len(layers) > 9:
            return False
        for L in layers:
            x, dy, dw = L['x'], L['dy'], L['dw']
            if tuple(dw.shape[:2]) != (5, 5) or dw.shape[3] != 4 or dw.shape[2] not in (1, 4) or x.shape[:3] != dy.shape[:3]:
                return False
            if dy.data_ptr() % 16 or dy.stride(0) % 4 or dy.stride(1) % 4 or dy.stride(3) != 1 or dy.stride(2) != 4:
                return False
            if dw.shape[2] == 4 and (x.data_ptr() % 16 or x.stride(0) % 4 or x.stride(1) % 4 or x.stride(3) != 1 or x.stride(2) != 4):
                return False
            if dw.shape[2] == 1 and x.stride(2) != 1:
                return False
        N, T, F, _ = layers[0]['dy'].shape
        return 2 * ((20 * (F + 4)) + 16 * F) * 16 <= 200 * 1024

    def conv2d_wgrad_multi(self, layers, scale=1.0, accumulate=False, bias_samples=None):
        """Weight (and bias) gradients of a stack of 5x5 'same' Conv2D layers in one launch (conv2d_wgrad44.cu).  layers: dicts with
        x [N,T,F,cin], dy [N,T,F,4], dw [5,5,cin,4], optional dbias [4] (sum over the first bias_samples samples)."""
        N, T, F, _ = layers[0]['dy'].shape
        n = len(layers)
        wkey = (self._tag, self._lane, 'multi', n)
        ws = self._wg.get(wkey)
        if ws is None:
            ws = self._wg[wkey] = self.empty(int(self.lib.pctts_conv2d_wgrad_multi_workspace(n)))
            self._note(self._wg, wkey)
        arr = (WgradLayerStruct * n)()
        for i, L in enumerate(layers):
            x, dy, dw, db = L['x'], L['dy'], L['dw'], L.get('dbias')
            assert tuple(dy.shape) == (N, T, F, 4) and tuple(x.shape[:3]) == (N, T, F) and dw.is_contiguous()
            assert db is None or (db.numel() == 4 and db.is_contiguous())
            arr[i] = WgradLayerStruct(x.data_ptr(), dy.data_ptr(), dw.data_ptr(), _ptr(db), x.stride(0), x.stride(1), dy.stride(0),
                                      dy.stride(1), int(dw.shape[2]))
        nb = N if bias_samples is None else int(bias_samples)
        self._ck(self.lib.pctts_conv2d_wgrad_multi(ctypes.cast(arr, ctypes.c_void_p), n, nb, ws.data_ptr(), N, T, F, float(scale),
                                                   int(accumulate), self._stream()), 'pctts_conv2d_wgrad_multi')

    # -- batch normalisation ---------------------------------------------------------------------------------
    def bn_stats(self, x, final=True):
        """-> ReduceWorkspace whose .sums = [sum x, sum x^2] over the rows of the contiguous [rows, C] tensor x.
        final=False (single replica, nothing to all-reduce): first stage only; bn_finalize() runs the second stage."""
        rows, C = x.shape
        assert x.is_contiguous()
        ws = self.reduce_ws(rows, C)
        if final:
            self._ck(self.lib.pctts_bn_stats(x.data_ptr(), rows, C, ws.ptr, self._stream()), 'pctts_bn_stats')
            ws.pending = 0
        else:
            self._ck(self.lib.pctts_bn_stats_partial(x.data_ptr(), rows, C, ws.ptr, self._stream()), 'pctts_bn_stats_partial')
            ws.pending = self.lib.pctts_colreduce_nblk(rows, C)
        return ws

    def bn_finalize(self, ws, count, eps, momentum, mean, invstd, moving_mean=None, moving_var=None):
        if getattr(ws, 'pending', 0):
            self._ck(self.lib.pctts_bn_finalize_partials(ws.ptr, ws.rows, ws.C, ws.pending, float(count), float(eps), float(momentum),
                                                         mean.data_ptr(), invstd.data_ptr(), _ptr(moving_mean), _ptr(moving_var),
                                                         self._stream()), 'pctts_bn_finalize_partials')
            ws.pending = 0
            return
        self._ck(self.lib.pctts_bn_finalize(ws.ptr, ws.rows, ws.C, float(count), float(eps), float(momentum), mean.data_ptr(),
                                            invstd.data_ptr(), _ptr(moving_mean), _ptr(moving_var), self._stream()),
                 'pctts_bn_finalize')

    def _plane_target(self, y):
        """Where the kernel that produces the contiguous [rows, C] matrix y may write y's bf16 operand planes itself (saving the
        pctts_split_bf16 launch in front of the consuming GEMM): (shadow, mark_fresh) or (None, None)."""
        if not (self.nterms > 0 and self._tracking and self.fuse_planes and y.dim() == 2 and y.is_contiguous()
                and y.shape[1] % 8 == 0 and y.shape[1] <= 1024 and y.data_ptr() % 16 == 0):
            return None, None
        par, r0 = self._row_parent(y)
        if par is not None:
            r1 = r0 + y.shape[0]
            return ShadowRows(par.shadow, r0, y.shape[0]), (lambda: par.add(r0, r1))
        key = (y.data_ptr(), tuple(y.shape), tuple(y.stride()))
        if self._is_persistent(key[0]):
            return None, None
        sh = self._shadows.get(key)
        if sh is None:
            sh = self._shadows[key] = Shadow(self.device, y.shape[0], y.shape[1], PLANES[self.nterms])
            self._note(self._shadows, key)
        ext = _extent(y)

        def mark():
            self._fresh[key] = ext
        return sh, mark

    def bn_apply(self, x, y, mean, invstd_or_var, gamma, beta, eps, use_var, lrelu, planes=False):
        """planes=True: y feeds a tensor-core GEMM -- its bf16 operand planes are written by the same kernel where possible."""
        rows, C = x.shape
        assert x.is_contiguous() and y.is_contiguous()
        self._wrote(y)
        sh, mark = self._plane_target(y) if planes else (None, None)
        if sh is not None:
            rc = self.lib.pctts_bn_apply_planes(x.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd_or_var.data_ptr(),
                                                gamma.data_ptr(), beta.data_ptr(), rows, C, float(eps), int(use_var), int(lrelu),
                                                sh.ptr, sh.ld, sh.part_stride, sh.parts, self._stream())
            if rc != -2:
                self._ck(rc, 'pctts_bn_apply_planes')
                mark()
                return
        self._ck(self.lib.pctts_bn_apply(x.data_ptr(), y.data_ptr(), mean.data_ptr(), invstd_or_var.data_ptr(),
                                         gamma.data_ptr(), beta.data_ptr(), rows, C, float(eps), int(use_var), int(lrelu),
                                         self._stream()), 'pctts_bn_apply')

    def bn_bwd_reduce(self, x, dy, mean, invstd, gamma, beta, lrelu, final=True):
        """final=False: first stage only (ws.pending partials), for bn_sync_bwd()."""
        rows, C = x.shape
        assert x.is_contiguous() and dy.is_contiguous()
        ws = self.reduce_ws(rows, C)
        fn = self.lib.pctts_bn_bwd_reduce if final else self.lib.pctts_bn_bwd_reduce_partial
        self._ck(fn(x.data_ptr(), dy.data_ptr(), mean.data_ptr(), invstd.data_ptr(), gamma.data_ptr(),
                    beta.data_ptr(), rows, C, int(lrelu), ws.ptr, self._stream()), 'pctts_bn_bwd_reduce')
        ws.pending = 0 if final else self.lib.pctts_colreduce_nblk(rows, C)
        return ws

    # -- BatchNorm second stage + replica exchange + finalize in ONE launch (csrc/peer.cu: bn_sync_kernel) --------------------
    bn_sync = True

    def _peer_args(self, sync):
        h = getattr(sync, 'bn_handle2' if self._lane == 2 else 'bn_handle', None)
        if h is None or sync.world == 1:
            return None, 0, 1, None
        bases, err = h
        return ctypes.cast(bases, ctypes.c_void_p), sync.rank, sync.world, err.data_ptr()

    def bn_sync_finalize(self, ws, count, eps, momentum, mean, invstd, moving_mean, moving_var, sync):
        assert getattr(ws, 'pending', 0) > 0
        bases, rank, world, err = self._peer_args(sync)
        self._ck(self.lib.pctts_bn_sync_finalize(ws.ptr, ws.sums.data_ptr(), ws.pending, ws.C, float(count), float(eps), float(momentum),
                                                 mean.data_ptr(), invstd.data_ptr(), _ptr(moving_mean), _ptr(moving_var), bases, rank,
                                                 world, err, self._stream()), 'pctts_bn_sync_finalize')
        ws.pending = 0

    def bn_sync_bwd(self, ws, dgamma, dbeta, scale, accumulate, sync):
        assert getattr(ws, 'pending', 0) > 0
        bases, rank, world, err = self._peer_args(sync)
        self._ck(self.lib.pctts_bn_sync_bwd(ws.ptr, ws.sums.data_ptr(), ws.pending, ws.C, _ptr(dgamma), _ptr(dbeta), float(scale),
                                            int(accumulate), bases, rank, world, err, self._stream()), 'pctts_bn_sync_bwd')
        ws.pending = 0

    def bn_bwd_apply(self, ws, x, dy, dx, mean, invstd, gamma, beta, count, lrelu, dgamma, dbeta, grad_scale=1.0,
                     accumulate=False, planes=False):
        """planes=True: dx feeds tensor-core GEMMs -- its bf16 operand planes are written by the same kernel where possible."""
        rows, C = x.shape
        assert dx.is_contiguous()
        self._wrote(dx)
        sh, mark = self._plane_target(dx) if (planes and dgamma is None and dbeta is None) else (None, None)
        if sh is not None:
            rc = self.lib.pctts_bn_bwd_apply_planes(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                                    gamma.data_ptr(), beta.data_ptr(), rows, C, float(count), int(lrelu), ws.ptr,
                                                    sh.ptr, sh.ld, sh.part_stride, sh.parts, self._stream())
            if rc != -2:
                self._ck(rc, 'pctts_bn_bwd_apply_planes')
                mark()
                return
        self._ck(self.lib.pctts_bn_bwd_apply(x.data_ptr(), dy.data_ptr(), dx.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                             gamma.data_ptr(), beta.data_ptr(), rows, C, float(count), int(lrelu),
                                             _ptr(dgamma), _ptr(dbeta), float(grad_scale), int(accumulate), ws.ptr,
                                             self._stream()),
                 'pctts_bn_bwd_apply')

    def colsum(self, x, out, accumulate=False):
        rows, C = x.shape
        assert x.is_contiguous() and out.numel() == C
        ws = self.reduce_ws(rows, C)
        self._ck(self.lib.pctts_colsum(x.data_ptr(), rows, C, out.data_ptr(), int(accumulate), ws.ptr, self._stream()),
                 'pctts_colsum')

    # -- LSTM ------------------------------------------------------------------------------------------------
    def lstm_fwd(self, zx, U, hout, gates=None, cstate=None, cluster_size=8, batch_chunk=2):
        """zx/U/gates/cstate: lists with one entry per direction (1 or 2); hout [B,T,ndir*H]."""
        ndir = len(zx)
        B, T, H4 = zx[0].shape
        H = H4 // 4
        g = gates if gates is not None else [None] * ndir
        c = cstate if cstate is not None else [None] * ndir
        z1, u1, g1, c1 = (zx[1], U[1], g[1], c[1]) if ndir == 2 else (None, None, None, None)
        for t in list(zx) + list(U):
            assert t.is_contiguous()
        assert hout.is_contiguous() and hout.shape[2] == ndir * H
        self._wrote(hout)
        self._ck(self.lib.pctts_lstm_fwd(zx[0].data_ptr(), _ptr(z1), U[0].data_ptr(), _ptr(u1), hout.data_ptr(), _ptr(g[0]),
                                         _ptr(g1), _ptr(c[0]), _ptr(c1), B, T, H, ndir, hout.shape[2], cluster_size,
                                         batch_chunk, self._stream()), 'pctts_lstm_fwd')

    def lstm_bwd(self, dhout, U, gates, cstate, dz, cluster_size=8, batch_chunk=2):
        ndir = len(U)
        B, T, H4 = gates[0].shape
        H = H4 // 4
        u1, g1, c1, d1 = (U[1], gates[1], cstate[1], dz[1]) if ndir == 2 else (None, None, None, None)
        assert dhout.is_contiguous()
        for t_ in dz:
            self._wrote(t_)
        self._ck(self.lib.pctts_lstm_bwd(dhout.data_ptr(), U[0].data_ptr(), _ptr(u1), gates[0].data_ptr(), _ptr(g1),
                                         cstate[0].data_ptr(), _ptr(c1), dz[0].data_ptr(), _ptr(d1), B, T, H, ndir,
                                         dhout.shape[2], cluster_size, batch_chunk, self._stream()), 'pctts_lstm_bwd')

    # -- WGAN-GP pieces --------------------------------------------------------------------------------------
    def critic_pack_inputs(self, real, fake, alpha, s0, S, off, skip_real=False):
        B, T, O = real.shape
        assert real.is_contiguous() and fake.is_contiguous() and s0.is_contiguous() and tuple(s0.shape) == (3 * B, T, S)
        self._wrote(s0)
        self._ck(self.lib.pctts_critic_pack_inputs(real.data_ptr(), fake.data_ptr(), alpha.data_ptr(), s0.data_ptr(), B, T, O, S,
                                                   off, int(skip_real), self._stream()), 'pctts_critic_pack_inputs')

    def fill_segments(self, p, n0, v0, n1=0, v1=0.0, n2=0, v2=0.0):
        assert p.is_contiguous() and p.numel() == n0 + n1 + n2
        self._wrote(p)
        self._ck(self.lib.pctts_fill_segments(p.data_ptr(), n0, float(v0), n1, float(v1), n2, float(v2), self._stream()),
                 'pctts_fill_segments')

    def sum_scaled(self, x, scale, losses, slot):
        assert x.is_contiguous()
        self._ck(self.lib.pctts_sum_scaled(x.data_ptr(), x.numel(), float(scale), losses.data_ptr(), slot, self._stream()),
                 'pctts_sum_scaled')

    def gp_norms(self, g, ss):
        B = g.shape[0]
        assert g.is_contiguous() and ss.numel() >= B
        self._ck(self.lib.pctts_gp_norms(g.data_ptr(), B, g.numel() // B, ss.data_ptr(), self._stream()), 'pctts_gp_norms')

    def gp_scale(self, g, ss, r0, lam, Bglobal, losses, slot):
        B = g.shape[0]
        assert g.is_contiguous() and r0.is_contiguous()
        self._wrote(r0)
        self._ck(self.lib.pctts_gp_scale(g.data_ptr(), ss.data_ptr(), r0.data_ptr(), B, g.numel() // B, float(lam), int(Bglobal),
                                         losses.data_ptr(), slot, self._stream()), 'pctts_gp_scale')

    def loss_combine(self, losses, out, a, ca, b=-1, cb=0.0, c=-1, cc=0.0):
        self._ck(self.lib.pctts_loss_combine(losses.data_ptr(), out, a, float(ca), b, float(cb), c, float(cc), self._stream()),
                 'pctts_loss_combine')

    def lse_loss_grad(self, pred, real, w, dpred, scale_loss, scale_grad, losses, slot):
        O = pred.shape[-1]
        rows = pred.numel() // O
        assert pred.is_contiguous() and real.is_contiguous() and (dpred is None or dpred.is_contiguous())
        if self._lse_ws is None:
            self._lse_ws = torch.zeros(self.lib.pctts_lse_workspace_bytes(), dtype=torch.uint8, device=self.device)
        self._wrote(dpred)
        self._ck(self.lib.pctts_lse_loss_grad(pred.data_ptr(), real.data_ptr(), _ptr(w), _ptr(dpred), rows, O, float(scale_loss),
                                              float(scale_grad), losses.data_ptr(), slot, self._lse_ws.data_ptr(),
                                              self._stream()), 'pctts_lse_loss_grad')

    def lse_loss_grad_cols(self, pred, real, w, dpred, c0, nc, scale_loss, scale_grad, losses, slot):
        """lse_loss_grad over the columns [c0, c0 + nc) only (loss sum -> losses[slot], gradient -> those columns of dpred)."""
        O = pred.shape[-1]
        rows = pred.numel() // O
        assert pred.is_contiguous() and real.is_contiguous() and (dpred is None or dpred.is_contiguous())
        ws = self._lse_ws_lane.get(self._lane)
        if ws is None:
            ws = self._lse_ws_lane[self._lane] = torch.zeros(self.lib.pctts_lse_workspace_bytes(), dtype=torch.uint8, device=self.device)
        if dpred is not None:
            self._wrote(dpred)
        self._ck(self.lib.pctts_lse_loss_grad_cols(pred.data_ptr(), real.data_ptr(), _ptr(w), _ptr(dpred), rows, O, int(c0), int(nc),
                                                   float(scale_loss), float(scale_grad), losses.data_ptr(), slot, ws.data_ptr(),
                                                   self._stream()), 'pctts_lse_loss_grad_cols')

    def adam(self, p, g, m, v, lr, beta1, beta2, eps, step):
        assert p.is_contiguous() and g.is_contiguous() and step.dtype == torch.int32
        self._wrote(p)
        self._ck(self.lib.pctts_adam(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), float(lr), float(beta1),
                                     float(beta2), float(eps), step.data_ptr(), self._stream()), 'pctts_adam')

    # -- small movers ----------------------------------------------------------------------------------------
    def add3(self, out, a, b, c):
        assert out.is_contiguous() and a.is_contiguous() and b.is_contiguous() and c.is_contiguous()
        self._wrote(out)
        self._ck(self.lib.pctts_add3(out.data_ptr(), a.data_ptr(), b.data_ptr(), c.data_ptr(), out.numel(), self._stream()), 'pctts_add3')

    def add(self, out, a, b):
        assert out.is_contiguous() and a.is_contiguous() and b.is_contiguous()
        self._wrote(out)
        self._ck(self.lib.pctts_add(out.data_ptr(), a.data_ptr(), b.data_ptr(), out.numel(), self._stream()), 'pctts_add')

    def sigmoid_bwd(self, dy, y, dz):
        rows, cols = dy.shape
        self._wrote(dz)
        self._ck(self.lib.pctts_sigmoid_bwd(dy.data_ptr(), dy.stride(0), y.data_ptr(), y.stride(0), dz.data_ptr(), dz.stride(0),
                                            rows, cols, self._stream()), 'pctts_sigmoid_bwd')

    def copy2d(self, dst, src, accumulate=False):
        rows, cols = dst.shape
        assert tuple(src.shape) == (rows, cols)
        self._wrote(dst)
        self._ck(self.lib.pctts_copy2d(dst.data_ptr(), dst.stride(0), src.data_ptr(), src.stride(0), rows, cols, int(accumulate),
                                       self._stream()), 'pctts_copy2d')

    def pad_time(self, dst, src, left, right):
        B, T, C = src.shape
        assert dst.is_contiguous() and src.is_contiguous() and tuple(dst.shape) == (B, T + left + right, C)
        self._wrote(dst)
        self._ck(self.lib.pctts_pad_time(dst.data_ptr(), src.data_ptr(), B, T, C, left, right, self._stream()), 'pctts_pad_time')

    def uniform(self, out, seed, counter):
        assert counter.dtype == torch.int64
        self._ck(self.lib.pctts_uniform(out.data_ptr(), out.numel(), int(seed), counter.data_ptr(), self._stream()),
                 'pctts_uniform')

    # -- generation tail -------------------------------------------------------------------------------------
    def denorm_split(self, cmp, mean, std, S, NM, out=None, f0=None, spec=None, nm=None):
        """cmp [T,O] * std + mean (None: no de-normalisation) -> out [T,O], f0 [T] = exp(col 0), spec [T,S], nm [T,NM] (each
        optional): modeltts.py:154-165 (denormalise) + the slicing / exp of vocoders.py:194-201."""
        T, O = cmp.shape
        assert cmp.is_contiguous() and (mean is None) == (std is None)
        for t, shape in ((out, (T, O)), (f0, (T,)), (spec, (T, S)), (nm, (T, NM))):
            assert t is None or (tuple(t.shape) == shape and t.is_contiguous()), (shape,)
        if mean is not None:
            assert mean.numel() == O and std.numel() == O and mean.is_contiguous() and std.is_contiguous()
        self._ck(self.lib.pctts_denorm_split(cmp.data_ptr(), _ptr(mean), _ptr(std), T, O, S, NM, _ptr(out), _ptr(f0), _ptr(spec),
                                             _ptr(nm), self._stream()), 'pctts_denorm_split')

    # -- PML band compression --------------------------------------------------------------------------------
    def pml_linbnd2fwbnd(self, x, idx, wgt, out, log_abs):
        T, nbins = x.shape
        K = out.shape[1]
        assert x.is_contiguous() and out.is_contiguous() and idx.dtype == torch.int32
        self._ck(self.lib.pctts_pml_linbnd2fwbnd(x.data_ptr(), idx.data_ptr(), wgt.data_ptr(), out.data_ptr(), T, nbins, K,
                                                 int(log_abs), self._stream()), 'pctts_pml_linbnd2fwbnd')

    def pml_fwbnd2linbnd(self, x, idx, wgt, out, smooth, do_exp):
        T, K = x.shape
        nbins = out.shape[1]
        assert x.is_contiguous() and out.is_contiguous() and idx.dtype == torch.int32
        self._ck(self.lib.pctts_pml_fwbnd2linbnd(x.data_ptr(), idx.data_ptr(), wgt.data_ptr(), out.data_ptr(), T, K, nbins,
                                                 int(smooth), int(do_exp), self._stream()), 'pctts_pml_fwbnd2linbnd')
