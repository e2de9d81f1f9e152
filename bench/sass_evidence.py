"""SASS evidence per kernel of libllmt_kernels.so (no GPU needed): counts of the Blackwell mnemonics that prove the
tcgen05 / TMEM / TMA paths and, for the fused compute+collective GEMMs, the system-scope peer loads / reductions in
the SAME function as the UTCHMMA tiles.

    python bench/sass_evidence.py [profiles/sass_evidence.md]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'llm_training_b200', '_C', 'libllmt_kernels.so')
PATTERNS = collections.OrderedDict([
    ('UTCHMMA (tcgen05.mma)', r'\bUTCHMMA'), ('LDTM/STTM (tcgen05.ld/st)', r'\b(LDTM|STTM)'),
    ('UTMALDG (TMA load)', r'\bUTMALDG'), ('UTMASTG (TMA store)', r'\bUTMASTG'),
    ('UTMAREDG (TMA reduce-add)', r'\bUTMAREDG'), ('UTCBAR (tcgen05.commit)', r'\bUTCBAR'),
    ('SYNCS (mbarrier)', r'\bSYNCS'), ('LDG .SYS (peer / system-scope load)', r'\bLDG\.[A-Z0-9.]*SYS'),
    ('STG .SYS / RED / ATOM', r'\b(STG\.[A-Z0-9.]*SYS|RED\.|REDG\.|ATOM|ATOMG)'),
    ('LDGMC / multimem (in-switch reduce, multicast store)', r'\b(LDGMC|STGMC|REDGMC|[A-Z]*\.MC\b)'),
    ('HMMA (legacy mma.sync)', r'\bHMMA'),
])


def demangle(names):
    out = subprocess.run(['c++filt'], input='\n'.join(names), capture_output=True, text=True).stdout.split('\n')
    return dict(zip(names, out))


def main() -> None:
    dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'profiles', 'sass_evidence.md')
    sass = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True, check=True).stdout
    funcs = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r'\s*Function : (\S+)', line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None:
            funcs[cur].append(line)
    names = demangle(list(funcs))
    rows = []
    for f, lines in funcs.items():
        text = '\n'.join(lines)
        counts = [len(re.findall(p, text)) for p in PATTERNS.values()]
        short = re.sub(r'\(.*', '', names[f]).replace('void ', '').replace('llmt::', '')
        t = re.search(r'gemm_bf16_kernel<(.*)>', names[f])
        if t:
            a = [x.strip() for x in t.group(1).split(',')]
            short = (f'gemm_bf16_kernel<cta_group={a[0]}, A_mn={a[1][-4:] == "true" or a[1] == "1"}, '
                     f'B_mn={a[2] == "true" or a[2] == "1"}, out={"fp32" if a[3] == "float" else "bf16"}, '
                     f'kDistA={a[4]}, kDistC={a[5]}>')
        rows.append((short, counts, len(lines)))
    with open(dst, 'w') as fh:
        fh.write('# SASS evidence (`cuobjdump -sass llm_training_b200/_C/libllmt_kernels.so`, sm_100a)\n\n')
        fh.write('Counts of instructions per kernel.  `kDistA=2`: all-gather -> GEMM (land once); `kDistA=3`: ZeRO-3 '
                 'weight gather -> GEMM; `kDistC=true`: GEMM -> reduce-scatter (TMA reduce-add into the owner); the '
                 'system-scope loads of the pull warps and the UTCHMMA tiles are in the SAME function.\n\n')
        fh.write('| kernel | ' + ' | '.join(PATTERNS) + ' | SASS lines |\n|---|' + '---|' * (len(PATTERNS) + 1) + '\n')
        for short, counts, n in sorted(rows):
            fh.write(f'| `{short}` | ' + ' | '.join(str(c) for c in counts) + f' | {n} |\n')
        # excerpt: the pull-warp loop of the land-once kernel next to its MMA issue
        for key, title in (('Li2ELb0ELb0E13__nv_bfloat16Li2ELb0E', 'all-gather -> GEMM (cta_group 2, kDistA=2)'),
                           ('Li2ELb0ELb0E13__nv_bfloat16Li3ELb0E', 'weight gather -> GEMM (cta_group 2, kDistA=3)'),
                           ('Li2ELb0ELb0E13__nv_bfloat16Li0ELb1E', 'GEMM -> reduce-scatter (cta_group 2, kDistC)')):
            f = next((f for f in funcs if key in f and 'gemm_bf16_kernel' in f), None)
            if f is None:
                continue
            picks = [l.strip() for l in funcs[f]
                     if re.search(r'UTCHMMA|UTMALDG|UTMAREDG|UTMASTG|LDG\.[A-Z0-9.]*SYS|RED\.|REDG|LDTM', l)]
            fh.write(f'\n## {title}\n\n`{f[:110]}…` — every tensor / TMA / peer instruction, in program order '
                     f'(first 40 of {len(picks)}):\n\n```\n' + '\n'.join(picks[:40]) + '\n```\n')
    print(f'wrote {dst}: {len(rows)} kernels')


if __name__ == '__main__':
    main()
