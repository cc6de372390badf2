"""tools/sass_census.py -- count the SASS mnemonics that prove which hardware paths each kernel of
libaccblas_b200.so uses (B200_PROFILING.md): UTCHMMA / UTCBAR / LDTM (tcgen05 MMA, commit, TMEM load), UTMALDG (TMA
tensor load), SYNCS (mbarrier), DMMA (FP64 tensor), HMMA (mma.sync TF32), plus registers from the cubin.
usage: python tools/sass_census.py > profiles/r01_sass_census.txt      (no GPU needed: cuobjdump on the built library)"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "accelerate-blas_b200", "lib", "libaccblas_b200.so")
WANT = ("UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMAPF", "SYNCS", "DMMA", "HMMA", "DFMA", "FFMA", "LDS", "LDG", "STG", "NANOSLEEP")

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
fn, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r"\(anonymous namespace\)::", "", fn)
        fn = re.sub(r"\(.*$", "", fn)[:110]
        counts[fn] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and fn:
        op = m.group(1)
        for w in WANT:
            if op.startswith(w):
                counts[fn][w] += 1
print("# SASS mnemonic census of accelerate-blas_b200/lib/libaccblas_b200.so (cuobjdump -sass, sm_100a); tools/sass_census.py")
print("# " + " ".join(f"{w:>8}" for w in WANT) + "  kernel")
for fn, c in counts.items():
    if not any(c[w] for w in ("UTCHMMA", "UTMALDG", "DMMA", "HMMA")) and "kernel" not in fn:
        continue
    print("  " + " ".join(f"{c[w]:>8}" for w in WANT) + "  " + fn)
