#!/usr/bin/env python
"""SASS evidence for profiles/: per shipped kernel an instruction histogram (cuobjdump -sass of the in-tree
libddp_b200.so, sm_100a), the mnemonics that matter called out (FFMA2/FMUL2 = packed f32x2, UBLKCP = TMA bulk copy,
SYNCS = mbarrier, LDG/STG.E.*.256 = 256-bit global access, multimem = NVSwitch), and for the AdamW kernel the
packed arithmetic of one group as an excerpt.   python tools/sass_report.py > profiles/r02_sass.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-inference-with-pytorch-and-litellm_b200", "libddp_b200.so")
out = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True).stdout
funcs, cur, name = {}, None, None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = m.group(1)
        cur = funcs.setdefault(name, [])
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
    if m and cur is not None:
        cur.append(m.group(1).strip())
demangled = subprocess.run(["c++filt"], input="\n".join(funcs), stdout=subprocess.PIPE, text=True).stdout.splitlines()
names = dict(zip(funcs, demangled))
WANT = [  # (label, regex on the demangled name)
    ("fused clip+AdamW, bf16, FastClip, TMA-staged 2 stages x 2 CTAs/SM (default, device table)", r"adamw_bulk_kernel<ddpb200::DevTable, 1, 1, 2, 2>"),
    ("fused clip+AdamW, bf16, FastClip, register kernel 128-bit x 4 CTAs/SM, packed math (host tables)", r"adamw_kernel<ddpb200::DevTable, 1, 1, 1, 1, 4, 16, true>"),
    ("fused clip+AdamW, fp32, FastClip, TMA-staged 2 stages x 2 CTAs/SM (default, device table)", r"adamw_bulk_kernel<ddpb200::DevTable, 0, 1, 2, 2>"),
    ("fused clip+AdamW, fp32, FastClip, register kernel 256-bit x 2 CTAs/SM, scalar IEEE arithmetic (host tables)", r"adamw_kernel<ddpb200::DevTable, 0, 0, 0, 1, 2, 32, false>"),
    ("pack bf16 -> bf16 (+scale)", r"stream_kernel<ddpb200::DevTable, 1, 1, 0, 1, true, false, false>"),
    ("L2 norm bf16 (sum of squares)", r"stream_kernel<ddpb200::DevTable, 1, 1, 1, 1, false, true, false>"),
    ("fused allreduce+norm, bf16, multicast (multimem), unroll 4", r"fused_allreduce_kernel<1, 1, 2, false, 4>"),
    ("fused allreduce+norm, bf16, peer mode, 8 ranks", r"fused_allreduce_kernel<1, 0, 8, false, 4>"),
    ("sharded AdamW + parameter all-gather, bf16, peer mode, 8 ranks, FastClip", r"sharded_adamw_kernel<1, 0, 8, 1>"),
    ("fold of the chunk partials", r"fold_partials_kernel"),
    ("norm finalize", r"norm_finalize_kernel"),
]
KEY = re.compile(r"^(FFMA2|FMUL2|FADD2|FFMA|FMUL|FADD|MUFU|UBLKCP|SYNCS|LDG|STG|LDS|STS|LDGMC|MULTIMEM|RED|ATOM|F2FP|FMNMX3|FMNMX|FCHK|CALL|BAR|HMMA|UTC|UTMA)")
print(__doc__.split("\n")[0])
print("library:", os.path.relpath(LIB, ROOT), "| arch:", ",".join(sorted(set(re.findall(r"arch = (sm_\w+)", out)))))
for label, rx in WANT:
    hit = [f for f in funcs if re.search(rx, names[f])]
    if not hit:
        print(f"\n## {label}\n  (no kernel matches {rx})")
        continue
    f = hit[0]
    ins = funcs[f]
    ops = collections.Counter()
    full = collections.Counter()
    for i in ins:
        i = re.sub(r"^@!?U?P\d+\s+", "", i)
        op = i.split()[0]
        ops[op.split(".")[0]] += 1
        if KEY.match(op):
            full[op] += 1
    print(f"\n## {label}\n   {names[f][:160]}\n   {len(ins)} instructions")
    print("   histogram: " + ", ".join(f"{k} {v}" for k, v in ops.most_common(24)))
    print("   of note:   " + ", ".join(f"{k} {v}" for k, v in sorted(full.items()) if re.match(r"(FFMA2|FMUL2|FADD2|UBLKCP|SYNCS|LDG|STG|LDGMC|MUFU|FCHK|CALL|FMNMX3|F2FP|RED|ATOM)", k)))
    if "adamw_bulk_kernel" in names[f]:
        start = next((k for k, i in enumerate(ins) if i.startswith("MUFU.RSQ")), None)
        if start is not None:
            print("   excerpt (written-out sqrt -> division by sqrt(bc2) -> + eps -> written-out quotient, packed f32x2):")
            for i in ins[max(0, start - 4):start + 34]:
                print("      " + i)
        blk = [i for i in ins if i.startswith(("UBLKCP", "SYNCS"))][:8]
        print("   producer / barrier instructions: " + " | ".join(blk))
