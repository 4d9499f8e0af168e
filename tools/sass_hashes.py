#!/usr/bin/env python
"""Per-kernel hash of the SASS in libadvdiff.so (addresses stripped).  Used to show that a change which only ADDS
code paths leaves every already-validated kernel bit-identical:
    python tools/sass_hashes.py > /tmp/before.json;  <edit, rebuild>;  python tools/sass_hashes.py /tmp/before.json"""
import hashlib
import json
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "multigpu_advectiondiffusion_b200", "libadvdiff.so")


def hashes():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs, cur = {}, None
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur and "/*" in ln:
            funcs[cur].append(re.sub(r"/\*[0-9a-f]{4,}\*/", "", ln).strip())
    return {k: hashlib.md5("\n".join(v).encode()).hexdigest() for k, v in funcs.items()}


def family_hash(hashes_by_kernel, family):
    """one hash for the instantiations of a kernel template whose mangled names match the regular expression `family`
    (the ones a capture actually profiled, e.g. r"burgers3d_stage_pairIdLi0ELi1ELi[123]ELb0E": double, WENO5-JS,
    multi-GPU semantics, stages 1-3, no peer store) -- adding other instantiations to the library leaves it alone"""
    items = sorted((k, v) for k, v in hashes_by_kernel.items() if re.search(family, k))
    return hashlib.md5(json.dumps(items).encode()).hexdigest() if items else None


if __name__ == "__main__":
    new = hashes()
    if len(sys.argv) > 1:
        old = json.load(open(sys.argv[1]))
        print("kernels: %d -> %d" % (len(old), len(new)))
        print("changed (same name, different SASS):", [k for k in old if k in new and new[k] != old[k]])
        print("old kernels whose SASS exists under no name any more:", [k for k, v in old.items() if v not in set(new.values())])
        print("added:", [k for k in new if k not in old])
    else:
        print(json.dumps(new, indent=0))
