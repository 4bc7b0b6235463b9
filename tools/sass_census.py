#!/usr/bin/env python
"""profiles/sass_rNN.txt: instruction census per kernel of the shipped library (cuobjdump -sass) plus an excerpt of the
row / column pass around their TMA, mbarrier and 256-bit store instructions.  usage: python tools/sass_census.py > profiles/sass_r02.txt"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "addingdisparityfiltering_b200", "libdwls_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
pat = {"UTMALDG": r"UTMALDG", "UBLKCP": r"UBLKCP", "SYNCS": r"SYNCS", "STG.256": r"STG\.E\.(?:\w+\.)*256", "LDGSTS": r"LDGSTS", "SHFL": r"SHFL",
       "MUFU.RCP": r"MUFU\.RCP", "BAR": r"\bBAR\.", "VABSDIFF4": r"VABSDIFF4", "IDP.4A": r"IDP\.4A"}
funcs = re.split(r"\n\s*Function : ", txt)
def demangle(n):
    try:
        return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip() or n
    except Exception:
        return n
print("# SASS census of addingdisparityfiltering_b200/libdwls_b200.so (cuobjdump -sass, sm_100a)")
print("# totals: " + ", ".join("%s %d" % (k, len(re.findall(v, txt))) for k, v in pat.items()))
print("# UTMALDG = cp.async.bulk.tensor (TMA tile load), UBLKCP = cp.async.bulk (TMA 1-D bulk copy), SYNCS = mbarrier ops, STG.*.256 = st.global.v8.f32")
print("# kernel | " + " | ".join(pat))
for f in funcs[1:]:
    name = f.split("\n", 1)[0].strip()
    cnt = {k: len(re.findall(v, f)) for k, v in pat.items()}
    if cnt["UTMALDG"] or cnt["UBLKCP"] or cnt["STG.256"] or cnt["LDGSTS"]:
        print(re.sub(r"\(.*", "", demangle(name))[:110] + " | " + " | ".join(str(cnt[k]) for k in pat))
for want, title in (("hsolve_kernelILi2ELi256ELi8ELb0E", "row pass hsolve_kernel<2,256,8>"), ("vsolve_kernelILi9ELi2ELi8ELi0ELi1024ELb1ELb0E", "column pass vsolve_kernel<9,2,8,FLOAT,1024,WB>")):
    for f in funcs[1:]:
        if want in f.split("\n", 1)[0]:
            lines = [l for l in f.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
            keep = [i for i, l in enumerate(lines) if re.search(r"UTMALDG|UBLKCP|SYNCS|STG\.E\.(?:\w+\.)*256|MUFU\.RCP|SHFL\.(UP|DOWN)", l)]
            print("\n## excerpt: %s (%d SASS instructions; lines with TMA / mbarrier / 256-bit store / first RCP and shuffle)" % (title, len(lines)))
            shown, nrcp, nshf = set(), 0, 0
            for i in keep:
                l = lines[i]
                if "MUFU" in l:
                    nrcp += 1
                    if nrcp > 2: continue
                if "SHFL" in l:
                    nshf += 1
                    if nshf > 2: continue
                print(re.sub(r"\s+", " ", l.split("/*")[1].split("*/")[0] + ": " + l.split("*/", 1)[1].split("/*")[0]).strip())
            break
