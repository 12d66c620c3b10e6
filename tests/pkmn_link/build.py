"""Build tests/pkmn_link/libpkmn_link_tu.so: link_tu.c compiled against the reference's own pkmn.h (in place, from
/root/reference) and linked against oak_b200/liboakb200.so. Test infrastructure; only possible where the reference tree
exists (this container) — the built .so travels to the GPU box with the snapshot."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_INCLUDE = "/root/reference/cpp/extern/engine/src/include"
SO = os.path.join(HERE, "libpkmn_link_tu.so")


def build(force=False):
    """-> path of the built library, or None when the reference header is not available and nothing was built before."""
    src = os.path.join(HERE, "link_tu.c")
    lib = os.path.join(ROOT, "oak_b200", "liboakb200.so")
    if not os.path.exists(os.path.join(REF_INCLUDE, "pkmn.h")):
        return SO if os.path.exists(SO) else None
    if not force and os.path.exists(SO) and os.path.getmtime(SO) >= max(os.path.getmtime(src), os.path.getmtime(lib)):
        return SO
    # -Wl,--no-undefined: every pkmn.h name the unit references must be found in liboakb200.so at LINK time
    subprocess.check_call(["gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-Werror", "-fPIC", "-shared", "-I", REF_INCLUDE, "-o", SO, src,
                           "-L", os.path.dirname(lib), "-loakb200", "-Wl,--no-undefined", "-Wl,-rpath,$ORIGIN/../../oak_b200"])
    return SO


if __name__ == "__main__":
    print(build(force=True))
