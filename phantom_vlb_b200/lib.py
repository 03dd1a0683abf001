"""ctypes binding of libvlb_b200.so.  include/vlb_b200.h is the single source of truth: prototypes are parsed
from it at load time, so the Python side cannot drift from the C ABI.  There is no fallback: if the shared
library is missing or a call fails, a RuntimeError is raised."""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VLB_LIB_PATH") or os.path.join(HERE, "libvlb_b200.so")     # override: A/B builds during tuning
HEADER = os.path.join(os.path.dirname(HERE), "include", "vlb_b200.h")

_CTYPES = {
    "int": ctypes.c_int, "float": ctypes.c_float, "double": ctypes.c_double,
    "long long": ctypes.c_longlong, "unsigned long long": ctypes.c_ulonglong, "size_t": ctypes.c_size_t,
}
_lib = None
_protos = None


def parse_header(path=HEADER):
    """-> {name: (restype_str, [(type_str, arg_name), ...])} for every prototype in the header."""
    txt = open(path).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//[^\n]*", "", txt)
    protos = {}
    for m in re.finditer(r"(const char\*|int|void)\s+(vlb_\w+)\s*\(([^)]*)\)\s*;", txt):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        al = []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                mm = re.match(r"(.*?)(\w+)$", a)
                ty = mm.group(1).strip()
                al.append((ty, mm.group(2)))
        protos[name] = (ret, al)
    return protos


def _ctype(ty):
    if "*" in ty:
        return ctypes.c_void_p
    ty = ty.replace("const ", "").strip()
    return _CTYPES[ty]


def protos():
    global _protos
    if _protos is None:
        _protos = parse_header()
    return _protos


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libvlb_b200.so not built (%s); run `python -m phantom_vlb_b200.build`. "
                               "There is no fallback path." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        for name, (ret, args) in protos().items():
            fn = getattr(L, name)          # AttributeError if the symbol is not exported
            fn.argtypes = [_ctype(t) for t, _ in args]
            fn.restype = {"int": ctypes.c_int, "void": None, "const char*": ctypes.c_char_p}[ret]
        _lib = L
    return _lib


def last_error():
    return lib().vlb_last_error_string().decode("utf-8", "replace")


_prof = None       # when a list: (name, start_event, end_event, args) per call (tools/breakdown.py, bench.py roofline pass)


def profile_begin():
    global _prof
    _prof = []


def profile_end():
    """-> [(name, milliseconds, args)] for every call since profile_begin() (synchronises the device)."""
    global _prof
    import torch
    torch.cuda.synchronize()
    rows, _prof = _prof, None
    return [(n, e0.elapsed_time(e1), a) for n, e0, e1, a in rows]


def call(name, *args):
    if _prof is not None and name != "vlb_check_device":
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = getattr(lib(), name)(*args)
        e1.record()
        _prof.append((name, e0, e1, args))
    else:
        rc = getattr(lib(), name)(*args)
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (name, rc, last_error()))


def ptr(t):
    """Raw device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def cur_stream():
    import torch
    return torch.cuda.current_stream().cuda_stream
