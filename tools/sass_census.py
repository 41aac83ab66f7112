"""Opcode census of the built library: per kernel, how many tcgen05 / TMA / TMEM instructions its SASS holds.
    python tools/sass_census.py > profiles/rNN_sass_census.txt
UTCHMMA = tcgen05.mma (kind::f16), UTCBAR = tcgen05.commit, LDTM = tcgen05.ld, UTMALDG / UTMASTG = cp.async.bulk.tensor
load / store (TMA), UBLKCP = cp.async.bulk (1-D bulk copy), UTCATOMSWS = tcgen05.alloc/dealloc handshakes,
SYNCS = mbarrier ops, HMMA/IMMA = warp-level mma.sync (none expected)."""
import collections
import hashlib
import re
import subprocess
import sys
from pathlib import Path

SO = Path(__file__).resolve().parent.parent / "celebrity-image-denoiser_b200" / "csrc" / "libcid_b200.so"
OPS = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "SYNCS", "HMMA", "IMMA", "FFMA", "MUFU"]


def main():
    so = Path(sys.argv[1]) if len(sys.argv) > 1 else SO
    sass = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True).stdout
    demangle = {}
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
            kernels[cur]["_total"] += 1
    names = list(kernels)
    if names:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        if len(out) == len(names):
            demangle = dict(zip(names, out))
    print(f"library {so.name} sha256 {hashlib.sha256(so.read_bytes()).hexdigest()[:16]}  ({len(kernels)} kernels, sm_100a SASS)")
    print("kernel".ljust(86) + " ".join(o.rjust(8) for o in OPS) + "   total")
    tot = collections.Counter()
    for k, c in sorted(kernels.items(), key=lambda kv: demangle.get(kv[0], kv[0])):
        name = demangle.get(k, k)
        name = re.sub(r"^void ", "", name).replace("cid::", "")
        name = re.sub(r"\((?:int|bool|unsigned int)\)", "", name)
        name = re.sub(r"\(.*$", "", name)
        print(name[:85].ljust(86) + " ".join(str(c.get(o, 0)).rjust(8) for o in OPS) + f"  {c['_total']:6d}")
        tot.update(c)
    print("ALL".ljust(86) + " ".join(str(tot.get(o, 0)).rjust(8) for o in OPS) + f"  {tot['_total']:6d}")
    tc = [k for k, c in kernels.items() if c.get("UTCHMMA")]
    print(f"\nkernels with tcgen05.mma (UTCHMMA): {len(tc)}; with TMA tensor loads (UTMALDG): "
          f"{sum(1 for c in kernels.values() if c.get('UTMALDG'))}; warp-level HMMA/IMMA anywhere: "
          f"{sum(c.get('HMMA', 0) + c.get('IMMA', 0) for c in kernels.values())}")


if __name__ == "__main__":
    main()
