"""SASS instruction counts of lumen_b200/libb200clip.so (the mnemonics B200_PROFILING.md lists as proof of tcgen05 / TMEM / TMA use):
    python profiles/sass_summary.py r2   ->  profiles/sass_summary_r2.txt"""
import collections, os, re, subprocess, sys
tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
here = os.path.dirname(os.path.abspath(__file__))
lib = os.path.join(here, "..", "lumen_b200", "libb200clip.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
PAT = [("UTCHMMA (tcgen05.mma)", r"\bUTCHMMA"), ("  of which .2CTA", r"\bUTCHMMA\.2CTA"), ("LDTM (tcgen05.ld)", r"\bLDTM"), ("STTM (tcgen05.st)", r"\bSTTM"),
       ("UTMALDG (TMA load)", r"\bUTMALDG"), ("UTMASTG (TMA store)", r"\bUTMASTG"), ("UTMAREDG (TMA reduce)", r"\bUTMAREDG"),
       ("UTMACCTL.PF (tensor-map prefetch)", r"\bUTMACCTL\.PF"), ("UTCBAR (tcgen05.commit)", r"\bUTCBAR"), ("legacy HMMA (mma.sync)", r"\bHMMA"),
       ("IDP.4A (dp4a)", r"\bIDP\.4A"), ("LDGSTS (cp.async)", r"\bLDGSTS")]
total = collections.Counter()
per = []
for i, body in enumerate(re.split(r"Function : \S+", sass)[1:]):
    c = collections.Counter({label: len(re.findall(rx, body)) for label, rx in PAT})
    total.update(c)
    if any(c.values()):
        nm = names[i].split("(")[0]
        per.append((nm, c))
L = [f"SASS summary of lumen_b200/libb200clip.so (cuobjdump -sass, sm_100a), round {tag[1:]}", "", "whole library:"]
for label, _ in PAT:
    L.append(f"  {label:<28} {total[label]}")
L += ["", "per kernel (only kernels that use one of the instructions above):"]
for nm, c in per:
    L.append(f"  {nm}")
    L.append("      " + ", ".join(f"{k.strip()} {v}" for k, v in c.items() if v and not k.startswith("  ")))
open(os.path.join(here, f"sass_summary_{tag}.txt"), "w").write("\n".join(L) + "\n")
print("\n".join(L[:16]))
