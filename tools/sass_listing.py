"""SASS evidence for profiles/: `cuobjdump -sass` of the in-tree librt_b200.so, one file per kernel the rooflines quote
(k_shade, k_trace_flat<2> = Cornell, k_traverse_p<10,false> = pawn, k_traverse_p<31,false> = demo2, k_reduce_multimem) and
a summary of the mnemonics that back the design claims (bulk copies + mbarrier, 256-bit read-only loads, multimem reduce,
no tensor-core / TMEM instructions: nothing on this path is a contraction).  No GPU needed.
usage: python tools/sass_listing.py profiles/r02_sass"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
prefix = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass")
sass = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "raytrace_b200", "librt_b200.so")], capture_output=True, text=True).stdout
blocks, name, cur = {}, None, []
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if name:
            blocks[name] = cur
        name, cur = m.group(1), []
    elif name:
        cur.append(line)
if name:
    blocks[name] = cur
demangled = subprocess.run(["c++filt"], input="\n".join(blocks), capture_output=True, text=True).stdout.splitlines()
by_name = dict(zip(demangled, blocks.values()))
WANT = {"k_shade": "rtb::k_shade<false, false>(", "k_trace_flat_2": "rtb::k_trace_flat<2u>(", "k_traverse_p_10": "rtb::k_traverse_p<10u, false>(",
        "k_traverse_p_31": "rtb::k_traverse_p<31u, false>(", "k_reduce_multimem": "rtb::k_reduce_multimem(", "k_raygen": "rtb::k_raygen(",
        "k_shade_bound": "rtb::k_shade<true, false>(", "k_build_tree": "rtb::gpubuild::k_build_tree("}
KEY = ["UBLKCP", "SYNCS", "LDG.E.ENL2.256", "LDG.E.128", "LDGMC", "ATOMG", "RED.E", "ATOMS", "STG.E.128", "STG.E.EF", "LDS", "STS", "FFMA", "MUFU",
       "SHFL", "VOTE", "BAR.SYNC", "BSSY", "HMMA", "UTCMMA", "UTCHMMA", "TCGEN", "LDTM", "DFMA", "DADD", "DMUL"]
rows = []
for tag, needle in WANT.items():
    hit = [k for k in by_name if needle in k]
    if not hit:
        rows.append(f"{tag}: NOT FOUND")
        continue
    body = by_name[hit[0]]
    ins = [re.sub(r"/\*[0-9a-f]{4,}\*/", "", l).strip() for l in body if re.search(r"/\*[0-9a-f]{4,}\*/", l)]
    # listing without the hex encodings (address + instruction only): a third of the size, same evidence
    text = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l).rstrip() for l in body]
    text = [l for l in text if l.strip()]
    with open(f"{prefix}_{tag}.txt", "w") as f:
        f.write(f"// cuobjdump -sass raytrace_b200/librt_b200.so : {hit[0]}   (hex encodings stripped)\n" + "\n".join(text) + "\n")
    ops = collections.Counter()
    for l in ins:
        m = re.match(r"(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m:
            ops[m.group(1)] += 1
    counts = {k: sum(v for o, v in ops.items() if o.startswith(k)) for k in KEY}
    rows.append(f"{tag} ({hit[0][:90]}): {sum(ops.values())} instructions; " + ", ".join(f"{k} {v}" for k, v in counts.items() if v) +
                "; tensor-core / TMEM instructions: " + str(sum(counts[k] for k in ("HMMA", "UTCMMA", "UTCHMMA", "TCGEN", "LDTM"))))
allops = collections.Counter()
for l in sass.splitlines():
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
    if m:
        allops[m.group(1)] += 1
lib = {k: sum(v for o, v in allops.items() if o.startswith(k)) for k in ("UBLKCP", "SYNCS", "LDG.E.ENL2.256", "LDGMC", "ATOMG", "HMMA", "UTCMMA", "LDTM")}
with open(f"{prefix}_summary.txt", "w") as f:
    f.write("# tools/sass_listing.py: mnemonic counts per kernel (full listings: " + os.path.basename(prefix) + "_<kernel>.txt)\n")
    f.write("\n".join(rows) + "\n")
    f.write("whole library: " + ", ".join(f"{k} {v}" for k, v in lib.items()) + "\n")
print(open(f"{prefix}_summary.txt").read())
