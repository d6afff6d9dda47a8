"""Import the UNMODIFIED reference packages installed under baseline/_ref (see build_ref.py).

BASELINE INFRASTRUCTURE: used by bench.py's reference arm and by the GPU acceptance tests, never by the product
package.  Nothing here reads /root/reference at run time.

The reference's packages carry the same import names as this repository's drop-in packages
(`diff_gaussian_rasterization`, `diff_gaussian_rasterization_amr`).  Two ways to get at them:

* `load(pkg)` in a process that has not imported the drop-in of that name: baseline/_ref goes first on sys.path
  and the package is imported under its own name -- exactly what a user of the reference does.
* `load(pkg, alias=True)` next to the drop-in (same-process comparisons): the package's stock `__init__.py` is
  executed from its file under the module name `gsr_ref_<pkg>`; its `from . import _C` then resolves to the stock
  extension in the same directory.  No file is altered either way.
"""
from __future__ import annotations

import ctypes as C
import importlib
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
CALLERS = os.path.join(DEST, "callers")


def package_dir(pkg: str) -> str:
    return os.path.join(DEST, pkg)


def available(pkg: str = "diff_gaussian_rasterization") -> bool:
    d = package_dir(pkg)
    return os.path.exists(os.path.join(d, "__init__.py")) and any(f.startswith("_C") and f.endswith(".so") for f in os.listdir(d))


def callers_available() -> bool:
    return all(os.path.exists(os.path.join(CALLERS, c)) for c in
               ("gaussian_renderer/__init__.py", "gaussian_renderer_amr/__init__.py", "utils/sh_utils.py"))


def load(pkg: str = "diff_gaussian_rasterization", alias: bool = False):
    if not available(pkg):
        raise RuntimeError(f"{package_dir(pkg)} is missing: run `python baseline/build_ref.py` where /root/reference exists")
    if alias:
        name = "gsr_ref_" + pkg
        if name in sys.modules:
            return sys.modules[name]
        spec = importlib.util.spec_from_file_location(name, os.path.join(package_dir(pkg), "__init__.py"),
                                                      submodule_search_locations=[package_dir(pkg)])
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod
    have = sys.modules.get(pkg)
    if have is not None:
        if os.path.dirname(os.path.abspath(have.__file__)) == package_dir(pkg):
            return have
        raise RuntimeError(f"{pkg} is already imported from {have.__file__}; use load(pkg, alias=True)")
    if DEST in sys.path:
        sys.path.remove(DEST)
    sys.path.insert(0, DEST)               # wins over the repository root (which holds the drop-in of the same name)
    mod = importlib.import_module(pkg)
    got = os.path.dirname(os.path.abspath(mod.__file__))
    if got != package_dir(pkg):
        raise RuntimeError(f"imported {pkg} from {got}, expected the reference install {package_dir(pkg)}")
    return mod


def pageable_memory_access(device_index: int) -> bool:
    """cudaDevAttrPageableMemoryAccess (88).  The reference's AMR fork passes a HOST stack array to a kernel
    (diff-gaussian-rasterization-amr/cuda_rasterizer/rasterizer_impl.cu:610,644); through the stock binding that only
    works where the GPU can dereference pageable host memory.  Where it cannot, calling it would leave a sticky
    illegal-address error in the process, so callers check this first."""
    try:
        rt = C.CDLL("libcudart.so.12")
        v = C.c_int(0)
        if rt.cudaDeviceGetAttribute(C.byref(v), C.c_int(88), C.c_int(int(device_index))) != 0:
            return False
        return bool(v.value)
    except OSError:
        return False


class RegisteredStackWorker:
    """Runs callables on ONE worker thread whose whole stack is cudaHostRegister'ed (portable | mapped).

    Why: the reference's AMR fork hands a host STACK array to setAMRLevelsKernel
    (diff-gaussian-rasterization-amr/cuda_rasterizer/rasterizer_impl.cu:610,644).  On a device without pageable-memory
    access the stock binding therefore faults -- unless the calling thread's stack is registered with the driver, which
    makes that very host pointer valid on the device (unified addressing; needs
    cudaDevAttrCanUseHostPointerForRegisteredMem).  Nothing of the reference is modified: the same `_C` call is simply
    issued from this thread.  Test/baseline infrastructure only."""

    def __init__(self, device_index: int = 0, stack_bytes: int = 32 << 20):
        import queue
        import threading
        self._q, self._res = queue.Queue(), queue.Queue()
        self.error = None
        self._dev = int(device_index)
        old = threading.stack_size(stack_bytes)
        try:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        finally:
            threading.stack_size(old)
        ok = self._res.get()
        if ok is not True:
            self.error = str(ok)
            raise RuntimeError(f"cannot register the worker stack: {ok}")

    def _register_own_stack(self):
        libc = C.CDLL(None)
        rt = C.CDLL("libcudart.so.12")
        attr = C.create_string_buffer(128)                      # pthread_attr_t (56 bytes on x86-64 glibc)
        libc.pthread_self.restype = C.c_void_p
        if libc.pthread_getattr_np(C.c_void_p(libc.pthread_self()), attr) != 0:
            return "pthread_getattr_np failed"
        addr, size = C.c_void_p(0), C.c_size_t(0)
        if libc.pthread_attr_getstack(attr, C.byref(addr), C.byref(size)) != 0:
            return "pthread_attr_getstack failed"
        libc.pthread_attr_destroy(attr)
        can = C.c_int(0)
        rt.cudaSetDevice(C.c_int(self._dev))
        rt.cudaDeviceGetAttribute(C.byref(can), C.c_int(91), C.c_int(self._dev))     # cudaDevAttrCanUseHostPointerForRegisteredMem
        if not can.value:
            return "device cannot use host pointers of registered memory"
        # touch every page first (the stack is lazily mapped), then pin it: portable (1) | mapped (2)
        page = 4096
        buf = (C.c_char * size.value).from_address(addr.value)
        guard = 64 * 1024                                       # leave the guard region at the low end alone
        for off in range(guard, size.value - 256 * 1024, page):  # do not scribble over the live frames at the top
            buf[off] = b"\0"
        e = rt.cudaHostRegister(C.c_void_p(addr.value + guard), C.c_size_t(size.value - guard), C.c_uint(1 | 2))
        if e != 0:
            rt.cudaGetLastError()
            return f"cudaHostRegister(stack) failed with error {e}"
        return True

    def _run(self):
        try:
            self._res.put(self._register_own_stack())
        except Exception as ex:  # noqa: BLE001
            self._res.put(repr(ex))
            return
        while True:
            item = self._q.get()
            if item is None:
                return
            fn, args, kw = item
            try:
                self._res.put((True, fn(*args, **kw)))
            except BaseException as ex:  # noqa: BLE001
                self._res.put((False, ex))

    def call(self, fn, *args, **kw):
        self._q.put((fn, args, kw))
        ok, val = self._res.get()
        if not ok:
            raise val
        return val

    def close(self):
        self._q.put(None)
