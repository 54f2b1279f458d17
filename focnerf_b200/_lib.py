"""ctypes loader for libfoc_b200.so -- the only door between Python and the CUDA kernels.

There is NO fallback: if the shared library is missing or an entry point is absent this module raises, and so
does every operator built on it (BASELINE.json north_star: "no CPU fallback").  Prototypes are not restated
here; they are parsed out of include/foc_b200.h, so the header stays the single source of truth and a
signature drift between header and library shows up as a load error.
"""
import ctypes
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfoc_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "foc_b200.h")

_SCALARS = {
    "int": ctypes.c_int, "uint32_t": ctypes.c_uint32, "int32_t": ctypes.c_int32, "uint64_t": ctypes.c_uint64,
    "int64_t": ctypes.c_int64, "float": ctypes.c_float, "size_t": ctypes.c_size_t, "void": None,
}


def parse_header(path=HEADER_PATH):
    """-> {name: (restype, [argtypes])} for every `foc_*` prototype in the header."""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    src = re.sub(r"^\s*#.*$", " ", src, flags=re.M)
    protos = {}
    for m in re.finditer(r"\b(int|size_t|void)\s+(foc_\w+)\s*\(([^)]*)\)\s*;", src):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                    continue
                toks = [t for t in a.replace("const", " ").split() if t]
                ty = toks[0]
                if ty not in _SCALARS:
                    raise RuntimeError(f"foc_b200.h: unknown scalar type in `{a}` of {name}")
                argtypes.append(_SCALARS[ty])
        protos[name] = (_SCALARS[ret], argtypes)
    return protos


_lib = None
_protos = None


def lib():
    global _lib, _protos
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m focnerf_b200.build` (or __graft_entry__.build()). "
                "focnerf_b200 has no CPU / PyTorch fallback for its CUDA operators.")
        l = ctypes.CDLL(LIB_PATH)
        _protos = parse_header()
        for name, (ret, argtypes) in _protos.items():
            try:
                fn = getattr(l, name)
            except AttributeError as e:
                raise RuntimeError(f"libfoc_b200.so does not export `{name}` declared in include/foc_b200.h") from e
            fn.restype, fn.argtypes = ret, argtypes
        if l.foc_abi_version() != 1:
            raise RuntimeError("libfoc_b200.so ABI version mismatch")
        _lib = l
    return _lib


def declared_symbols():
    return sorted(parse_header().keys())


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def check(err, what):
    if err == 0:
        return
    if err == -1:
        raise RuntimeError(f"{what}: unsupported configuration for the sm_100a kernels")
    raise RuntimeError(f"{what}: CUDA error {err}")


# ---- device-side row count of a graph-replayed train step (include/foc_b200.h "device-side row counts").  The Trainer sets it around the
# step it captures; the march wrapper publishes into it and the grid / field / compositor wrappers pass it on.  None = every row counts.
_valid_rows = None


def set_valid_rows(t):
    """t: int32 [1] CUDA tensor (or None).  Returns the previous value."""
    global _valid_rows
    prev, _valid_rows = _valid_rows, t
    return prev


def valid_rows():
    return _valid_rows


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("focnerf_b200 operators need CUDA tensors (there is no CPU path)")


# ---------------------------------------------------------------------------------------------- per-entry-point timing
class KernelTimer:
    """Wraps selected `foc_*` entry points so every call is bracketed by CUDA events recorded on the launching
    (current) stream; also counts every call.  Used by bench.py for `roofline.achieved` and `gpu_launches`.
    Launch counts per entry point (kernels per call) are declared in LAUNCHES below."""

    # kernels launched by one call of each entry point (see the C ABI implementations in csrc/)
    LAUNCHES = {"foc_march_rays_train": 2, "foc_compact_alive": 2, "foc_ffmlp_backward": 2, "foc_sample_uniform": 2,
                "foc_grid_encode_backward": 1}

    def __init__(self, names=()):
        self.names, self.events, self.calls, self._orig = tuple(names), {}, {}, {}

    def __enter__(self):
        l = lib()
        for name in _protos:
            fn = getattr(l, name)
            self._orig[name] = fn
            self.calls[name] = 0
            timed = name in self.names
            if timed:
                self.events[name] = []

            def wrapper(*a, _fn=fn, _name=name, _timed=timed):
                self.calls[_name] += 1
                if _name == "foc_ffmlp_backward" and a[14] is None:      # grad_weights == NULL: no reduction kernel in this call
                    self.calls["_ffmlp_backward_no_finish"] = self.calls.get("_ffmlp_backward_no_finish", 0) + 1
                if not _timed:
                    return _fn(*a)
                s = torch.cuda.current_stream()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s)
                r = _fn(*a)
                e1.record(s)
                self.events[_name].append((e0, e1))
                return r
            setattr(l, name, wrapper)
        return self

    def __exit__(self, *exc):
        l = lib()
        for name, fn in self._orig.items():
            setattr(l, name, fn)
        return False

    def mean_ms(self, name):
        ev = self.events.get(name, [])
        return sum(a.elapsed_time(b) for a, b in ev) / max(1, len(ev))

    def kernel_launches(self):
        skip = ("foc_abi_version", "foc_allocate_splitk", "foc_free_splitk")
        total = sum(n * self.LAUNCHES.get(k, 1) for k, n in self.calls.items() if not k.endswith("_bytes") and k not in skip and not k.startswith("_"))
        return total - self.calls.get("_ffmlp_backward_no_finish", 0)


class GraphKernelTimer(KernelTimer):
    """KernelTimer for a CUDA-graph-replayed step: while a stream is being CAPTURED, every timed `foc_*` call is bracketed by two
    external event-record nodes (foc_timer_mark), so each replay of the graph stamps them and `collect()` (after a synchronize)
    reads the device time of every call as it ran inside the replayed step.  Calls made outside a capture are counted but not timed.

        with GraphKernelTimer(names) as kt:
            for batch in batches:
                kt.begin_step(); trainer.train_step(*batch); torch.cuda.synchronize(); kt.collect()
        kt.mean_ms(name), kt.calls_per_step(name)
    """
    SLOTS = 8192

    def __enter__(self):
        l = lib()
        h = ctypes.c_void_p()
        check(l.foc_timer_create(self.SLOTS, ctypes.byref(h)), "timer_create")
        self._h, self._next, self.pairs, self._mark = h, 0, [], 0
        self.sums, self.counts, self.steps = {}, {}, 0
        for name in _protos:
            fn = getattr(l, name)
            self._orig[name] = fn
            self.calls[name] = 0
            if name.startswith("foc_timer_"):
                continue
            timed = name in self.names

            def wrapper(*a, _fn=fn, _name=name, _timed=timed):
                self.calls[_name] += 1
                if _name == "foc_ffmlp_backward" and a[14] is None:
                    self.calls["_ffmlp_backward_no_finish"] = self.calls.get("_ffmlp_backward_no_finish", 0) + 1
                if not _timed or not torch.cuda.is_current_stream_capturing() or self._next + 2 > self.SLOTS:
                    return _fn(*a)
                st, k = stream(), self._next
                self._next += 2
                self._orig["foc_timer_mark"](self._h, k, st)
                r = _fn(*a)
                self._orig["foc_timer_mark"](self._h, k + 1, st)
                self.pairs.append((_name, k))
                return r
            setattr(l, name, wrapper)
        return self

    def __exit__(self, *exc):
        r = super().__exit__(*exc)
        lib().foc_timer_destroy(self._h)
        return r

    def begin_step(self):
        self._mark = len(self.pairs)

    def collect(self):
        """After the step and a synchronize.  A step that (re-)captured the graph invalidates the pairs of the previous graph."""
        if len(self.pairs) > self._mark and self._mark > 0:
            self.pairs = self.pairs[self._mark:]
            self.sums, self.counts, self.steps = {}, {}, 0
        if not self.pairs:
            return
        ms = ctypes.c_float()
        for name, k in self.pairs:
            if self._orig["foc_timer_elapsed"](self._h, k, k + 1, ctypes.byref(ms)) != 0:
                return                                   # not stamped yet (capture without a replay)
            self.sums[name] = self.sums.get(name, 0.0) + ms.value
            self.counts[name] = self.counts.get(name, 0) + 1
        self.steps += 1

    def mean_ms(self, name):
        return self.sums.get(name, 0.0) / max(1, self.counts.get(name, 0))

    def calls_per_step(self, name):
        return self.counts.get(name, 0) / max(1, self.steps)

    def timed_names(self):
        return [n for n in self.names if self.counts.get(n)]
