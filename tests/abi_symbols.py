"""Names of the functions include/spectavi_b200.h declares (parsed from the
header, so a new declaration is checked without touching the tests)."""
import os
import re

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "spectavi_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    names = []
    for m in re.finditer(r"^[A-Za-z_][\w \*]*?\b(\w+)\s*\(([^;{}]*)\)\s*;", src, flags=re.M):
        if "typedef" in m.group(0):
            continue
        names.append(m.group(1))
    return names
