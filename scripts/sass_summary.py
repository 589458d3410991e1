"""profiles/<tag>_sass_tcgen05.txt: Blackwell-native SASS instructions per kernel of librn_b200.so (counts + one example
line each), from `cuobjdump -sass` piped through c++filt.  usage: python scripts/sass_summary.py [tag=r02]"""
import os, re, subprocess, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
so = os.path.join(ROOT, "distributed-tensorflow-resnet_b200", "librn_b200.so")
sass = subprocess.run("cuobjdump -sass %s | c++filt" % so, shell=True, capture_output=True, text=True, check=True).stdout
PAT = re.compile(r"\b(UTCHMMA[.\w]*|UTCQMMA[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UTMAREDG[.\w]*|UTMAPF[.\w]*|UTCBAR[.\w]*|"
                 r"UTCATOMSWS[.\w]*|SYNCS[.\w]*|UCGABAR[.\w]*|MAPA[.\w]*|ACQBULK|UTMACMDFLUSH|UTMACCTL[.\w]*)")
cur, kernels = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.match(r"\s*Function : (.*)", line)
    if m:
        cur = m.group(1).strip()
        kernels[cur] = (collections.Counter(), {})
        continue
    if cur is None:
        continue
    m = PAT.search(line)
    if m and "/*" in line:
        cnt, ex = kernels[cur]
        cnt[m.group(1)] += 1
        key = m.group(1).split(".")[0] + (".2CTA" if ".2CTA" in m.group(1) else "")
        if key not in ex and re.match(r"UTCHMMA|UTMALDG|UTMASTG|UTMAREDG|LDTM|UTCBAR", key):
            ex[key] = re.sub(r"^\s*/\*[0-9a-f]+\*/\s*", "", line).rstrip()
out = ["# cuobjdump -sass librn_b200.so: Blackwell-native instructions per kernel (counts) and one example line each",
       "# UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM = tcgen05.ld, UTMALDG / UTMASTG / UTMAREDG = TMA tensor load /",
       "# store / reduce-add (.2CTA = cta_group::2 load signalling the pair leader), UTCBAR = tcgen05.commit (.MULTICAST for pairs)",
       "# regenerate: python scripts/sass_summary.py " + tag, ""]
for name, (cnt, ex) in kernels.items():
    if not any(k.startswith(("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM")) for k in cnt):
        continue
    out.append(name)
    out.append("   " + ", ".join("%s x%d" % kv for kv in sorted(cnt.items())))
    for k in sorted(ex):
        out.append("      " + ex[k])
    out.append("")
path = os.path.join(ROOT, "profiles", "%s_sass_tcgen05.txt" % tag)
open(path, "w").write("\n".join(out))
print("wrote", path, "(%d kernels)" % sum(1 for l in out if l.startswith(("void", "rn::"))))
