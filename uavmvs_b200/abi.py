"""The C ABI as declared in include/uavmvs_b200.h, parsed from the header itself, so the
"every declared symbol is exported" check cannot drift from the header."""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(_HERE, "..", "include", "uavmvs_b200.h")
SO = os.path.join(_HERE, "libuavmvs_b200.so")


def declared_symbols(header=HEADER):
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(uavmvs_[a-z0-9_]+)\s*\(", text)
    out = []
    for n in names:
        if n not in out:
            out.append(n)
    return out


def missing_symbols(so=SO):
    lib = ctypes.CDLL(so)
    return [n for n in declared_symbols() if not hasattr(lib, n)]
