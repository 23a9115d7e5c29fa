"""ctypes binding of the C-ABI library (include/explainn_b200.h).

The prototypes are read from the header itself so that the Python side can never
drift from the declared ABI.  There is no CPU or library fallback: if the shared
object is missing the import of any op fails loudly.
"""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(_HERE), "include", "explainn_b200.h")
LIB_PATH = os.path.join(_HERE, "lib", "libexplainn_b200.so")

_SCALARS = {
    "int": ctypes.c_int,
    "float": ctypes.c_float,
    "double": ctypes.c_double,
    "size_t": ctypes.c_size_t,
    "int64_t": ctypes.c_longlong,
    "long long": ctypes.c_longlong,
    "unsigned long long": ctypes.c_ulonglong,
    "xnn_stream_t": ctypes.c_void_p,
}


def parse_header(path=HEADER):
    """Returns {name: (restype, [argtypes], [argnames])} for every xnn_* prototype."""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    protos = {}
    for m in re.finditer(r"\b(int|size_t|const char\s*\*)\s+(xnn_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        ret, name, args = m.group(1), m.group(2), " ".join(m.group(3).split())
        restype = {"int": ctypes.c_int, "size_t": ctypes.c_size_t}.get(ret, ctypes.c_char_p)
        argtypes, argnames = [], []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                    argnames.append(a.split("*")[-1].strip())
                else:
                    toks = a.replace("const ", "").split()
                    ctype = " ".join(toks[:-1])
                    argtypes.append(_SCALARS[ctype])
                    argnames.append(toks[-1])
        protos[name] = (restype, argtypes, argnames)
    return protos


class XnnError(RuntimeError):
    pass


_lib = None


def load():
    """dlopen the library and attach prototypes; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise XnnError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  explainn_b200 has no CPU or library fallback."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes, _) in parse_header().items():
        fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(status, what=""):
    if status != 0:
        msg = load().xnn_last_error()
        raise XnnError(f"{what}: status {status}: {msg.decode() if msg else ''}")
