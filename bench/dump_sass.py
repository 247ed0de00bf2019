"""Regenerates profiles/sass/: `cuobjdump -sass` of libdrn_b200.so split per kernel, one file per
kernel (the largest instantiation of each template family is kept in full, the others are counted),
and SUMMARY.md with the mnemonic counts that identify the Blackwell paths
(UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA, multimem forms, REDG = red.global).
Runs on the CPU box (no GPU needed):  python bench/dump_sass.py"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed_resnet_tensorflow_b200", "_native", "libdrn_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
INTERESTING = re.compile(r"^(UTC\w+|LDTM|STTM|UTMA\w+|UBLKCP|SYNCS|REDG|RED|ATOMG|ATOMS|ATOM|LDGMC|STGMC|REDGMC|LDG|STG|"
                         r"F2FP|HMMA|ACQBULK|PREEXIT|UCGABAR|MEMBAR|CCTL|FENCE|ERRBAR|NANOSLEEP|MUFU)")


def demangle(names):
    # file-local kernels carry a "__nv_static_NN__<hash>_<file>__" prefix in front of the Itanium name
    stripped = [n[n.index("_ZN"):] if "_ZN" in n else n for n in names]
    out = subprocess.run(["c++filt"], input="\n".join(stripped), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def short(dem):
    m = re.search(r"::(\w+_kernel\w*)(<[^(]*>)?\(", dem) or re.search(r"(\w+_kernel\w*)(<[^(]*>)?", dem)
    if not m:
        return dem[:60], ""
    return m.group(1), (m.group(2) or "")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs, cur = {}, None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = funcs.setdefault(m.group(1), [])
        if cur is not None:
            cur.append(line)
    dem = demangle(list(funcs))
    fam = collections.defaultdict(list)
    for mangled, lines in funcs.items():
        if mangled.startswith(("__cuda_", "_INTERNAL_")) or "_kernel" not in mangled:
            continue                                   # compiler helpers (division slow paths, traps)
        base, targs = short(dem[mangled])
        ops = collections.Counter()
        n_inst = 0
        for ln in lines:
            m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", ln)
            if not m:
                continue
            n_inst += 1
            op = m.group(1)
            if INTERESTING.match(op):
                keep_full = op.startswith(("UTC", "UTMA", "LDGMC", "STGMC")) or ".MC" in op
                ops[op if keep_full else op.split(".")[0]] += 1
        fam[base].append((n_inst, targs, ops, lines))
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        if f.endswith(".sass"):
            os.remove(os.path.join(OUT, f))
    rows = []
    for base in sorted(fam):
        insts = sorted(fam[base], key=lambda t: -t[0])
        n_inst, targs, ops, lines = insts[0]
        with open(os.path.join(OUT, base + ".sass"), "w") as f:
            f.write(f"// {base}{targs}: largest of {len(insts)} instantiation(s), {n_inst} instructions\n")
            # keep the instruction text, drop the encoding words (halves the size)
            f.write("\n".join(re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", ln) for ln in lines
                              if not re.match(r"^\s*/\* 0x[0-9a-f]{16} \*/\s*$", ln)) + "\n")
        for n_inst_i, targs_i, ops_i, _ in insts:
            rows.append((base, targs_i, n_inst_i, ops_i))
    with open(os.path.join(OUT, "SUMMARY.md"), "w") as f:
        f.write("# SASS evidence (cuobjdump -sass of libdrn_b200.so, sm_100a; regenerate with bench/dump_sass.py)\n\n"
                "One `.sass` file per kernel (largest instantiation of a template family); every instantiation is\n"
                "listed below with its instruction count and the mnemonics that identify the hardware paths:\n"
                "`UTCHMMA` = tcgen05.mma (`.2CTA` = cta_group::2), `LDTM` = tcgen05.ld, `UTMALDG`/`UTMASTG` = TMA\n"
                "load/store (`.IM2COL`, `.2CTA`), `UTCBAR` = tcgen05.commit, `UTCATOMSWS` = TMEM allocation,\n"
                "`SYNCS` = mbarrier, `REDG` = red.global, `LDGMC` / `.MC` forms = multimem.ld_reduce / multimem.st\n"
                "through the NVSwitch, `ACQBULK`/`PREEXIT` = griddepcontrol, `F2FP` = packed bf16 conversion.\n\n"
                "| kernel | template arguments | instructions | mnemonics |\n|---|---|---|---|\n")
        for base, targs, n_inst, ops in rows:
            desc = ", ".join(f"{k}x{v}" for k, v in sorted(ops.items()))
            f.write(f"| `{base}` | `{targs[:90]}` | {n_inst} | {desc} |\n")
        f.write(f"\nHMMA (mma.sync) present anywhere: {any('HMMA' in ops for _, _, _, ops in rows)}\n")
    print(f"{len(funcs)} kernels, {len(fam)} families -> {OUT}")


if __name__ == "__main__":
    sys.exit(main())
