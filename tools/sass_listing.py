"""Writes trimmed SASS listings of the hot kernels into profiles/sass/ (BASELINE.json: "each kernel is shown in a
committed SASS listing"; B200_PROFILING.md: UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA,
LDGMC/... .ADD = multimem.ld_reduce, ACQBULK/PREEXIT = programmatic dependent launch).

Per kernel: (a) the instruction stream in full from a few instructions before the first tensor-core / TMA / multimem
instruction to a few after the last one (the main loop and the accumulator hand-over), (b) a "skeleton" of the whole
kernel: every tensor, TMA, TMEM, barrier, atomic, fence, system-scope and control-flow instruction with its offset.

    python tools/sass_listing.py            # reads pytorch_multiprocessing-distributed_b200/_lib/libpmdist.so
"""
from __future__ import annotations

import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'pytorch_multiprocessing-distributed_b200', '_lib', 'libpmdist.so')
OUT = os.path.join(ROOT, 'profiles', 'sass')

# output file -> substring that selects the (mangled) kernel name
KERNELS = {
    'conv_gemm_kernel_128_3_fprop': 'conv_gemm_kernelILi128ELi3ELb0ELi0ELb0EE',
    'conv_gemm_kernel_256_2_dgrad_krsc': 'conv_gemm_kernelILi256ELi2ELb1ELi0ELb0EE',
    'conv_gemm_kernel_128_3_dgrad_bn_reduce': 'conv_gemm_kernelILi128ELi3ELb1ELi1ELb0EE',
    'conv_gemm_kernel_128_3_dgrad_block_bn_reduce': 'conv_gemm_kernelILi128ELi3ELb1ELi2ELb0EE',
    'conv_gemm_kernel_128_3_fprop_cta_pair': 'conv_gemm_kernelILi128ELi3ELb0ELi0ELb1EE',
    'conv_wgrad_kernel_256_3_2': 'conv_wgrad_kernelILi256ELi3ELi2EE',
    'conv_wgrad_patch_kernel_4': 'conv_wgrad_patch_kernelILi4EE',
    'conv3x3_resident_kernel_fprop': 'conv3x3_resident_kernelILb0ELb0EE',
    'conv3x3_resident_kernel_dgrad_bn_reduce': 'conv3x3_resident_kernelILb1ELb1EE',
    'stem_fused_kernel_gather_4groups': 'stem_fused_kernelILb1ELi4EE',
    'allreduce_sgd_ll_kernel': 'allreduce_sgd_ll_kernel',
    'allreduce_sgd_kernel_nvls': 'allreduce_sgd_kernelILi3EE',
    'allreduce_sgd_kernel_twoshot': 'allreduce_sgd_kernelILi2EE',
    'small_allreduce_push_kernel': 'small_allreduce_push_kernel',
    'bn_finalize_push_kernel': 'bn_finalize_push_kernel',
    'augment_im2col_kernel_gather_philox': 'augment_im2col_kernelILb1EE',
}
CORE = re.compile(r'\b(UTC[A-Z]*MMA|UTCBAR|UTCATOMSWS|UTMALDG|UTMASTG|UTMACMDFLUSH|UTMACCTL|UBLKCP|LDTM|STTM|LDGMC|multimem)\b|'
                  r'\.SYS\b|SYNCS')
SKEL = re.compile(r'\b(UTC\w*|UTMA\w*|UBLKCP|LDTM|STTM|LDGMC|SYNCS|BAR|MEMBAR|FENCE|ERRBAR|ACQBULK|PREEXIT|EXIT|BRA|CALL|RET|'
                  r'RED|REDG|ATOM|ATOMG|ATOMS|ELECT|NANOSLEEP|WARPSYNC|CCTL|UCGABAR\w*|R2UR|S2UR)\b|\.SYS\b|\.STRONG\.')


def main() -> None:
    txt = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True, check=True).stdout
    funcs, cur = {}, None
    for ln in txt.splitlines():
        m = re.search(r'Function : (\S+)', ln)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur is None:
            continue
        m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*', ln)
        if m:
            funcs[cur].append((m.group(1), re.sub(r'\s+', ' ', m.group(2)).strip()))
    os.makedirs(OUT, exist_ok=True)
    index = []
    for out_name, key in KERNELS.items():
        match = [f for f in funcs if key in f]
        if not match:
            print('not found:', key, file=sys.stderr)
            continue
        name = match[0]
        ins = funcs[name]
        hot = [i for i, (_, t) in enumerate(ins) if CORE.search(t)]
        lo = max(0, (hot[0] if hot else 0) - 24)
        hi = min(len(ins), (hot[-1] if hot else 0) + 12)
        counts = {}
        for _, t in ins:
            op = t.split()[1] if t.startswith('@') and len(t.split()) > 1 else t.split()[0]
            op = op.split('.')[0]
            if re.match(r'UTC|UTMA|LDTM|STTM|LDGMC|ACQBULK|PREEXIT|SYNCS|UBLKCP|HMMA|HGMMA', op):
                counts[op] = counts.get(op, 0) + 1
        with open(os.path.join(OUT, out_name + '.sass'), 'w') as f:
            f.write(f'// {name}\n// sm_100a SASS from libpmdist.so (cuobjdump -sass), {len(ins)} instructions.\n')
            f.write('// mnemonic counts: ' + ', '.join(f'{k} {v}' for k, v in sorted(counts.items())) + '\n\n')
            f.write(f'// ---- (a) instructions {lo}..{hi} of {len(ins)}: first to last tensor/TMA/multimem/system-scope instruction\n')
            for off, t in ins[lo:hi]:
                f.write(f'/*{off}*/  {t}\n')
            f.write('\n// ---- (b) skeleton of the whole kernel\n')
            for off, t in ins:
                if SKEL.search(t):
                    f.write(f'/*{off}*/  {t}\n')
        index.append((out_name, name, len(ins), counts))
    with open(os.path.join(OUT, 'README.md'), 'w') as f:
        f.write('# SASS listings (sm_100a, from the in-tree libpmdist.so; regenerate with `python tools/sass_listing.py`)\n\n')
        f.write('| file | kernel | instructions | tensor / TMA / TMEM / PDL mnemonics |\n|---|---|---|---|\n')
        for out_name, name, n, counts in index:
            f.write(f'| `{out_name}.sass` | `{name[:90]}` | {n} | ' + ', '.join(f'{k} {v}' for k, v in sorted(counts.items())) + ' |\n')
        f.write('\nUTCHMMA = tcgen05.mma (bf16), LDTM = tcgen05.ld, UTMALDG / UTMASTG = cp.async.bulk.tensor load / store, '
                'UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, LDGMC...ADD = multimem.ld_reduce, ACQBULK / PREEXIT = '
                'griddepcontrol.wait / launch_dependents.  No HMMA / HGMMA (mma.sync / wgmma) anywhere.\n')
    print('wrote', len(index), 'listings to', OUT)


if __name__ == '__main__':
    main()
