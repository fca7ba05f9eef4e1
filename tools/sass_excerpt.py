"""SASS evidence for the tensor-core / TMA / multimem kernels (no GPU needed): for every kernel of the library that
uses them, the count of each Blackwell mnemonic and the first occurrence of each (with its address), from
`cuobjdump -sass vae_npvc_b200/lib/libvae_npvc_b200.so`.  Output: profiles/r02_sass_excerpt.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'vae_npvc_b200', 'lib', 'libvae_npvc_b200.so')
PAT = re.compile(r'\b(UTCHMMA[.\w]*|UTCBAR[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UTMAPF[.\w]*|UBLKCP[.\w]*|UBLKRED[.\w]*|LDTM[.\w]*|'
                 r'STTM[.\w]*|UTCATOMSWS[.\w]*|UTCCP[.\w]*|SYNCS[.\w]*|ELECT[.\w]*|LDGMC[.\w]*|STG[.\w]*MC[.\w]*|REDG?MC[.\w]*|'
                 r'ACQBULK|UCGABAR[.\w]*|FENCE[.\w]*|LDGSTS[.\w]*|FMNMX3[.\w]*|UTMACCTL[.\w]*|UTMACMDFLUSH)\b')


def main():
    sass = subprocess.run(['cuobjdump', '-sass', LIB], capture_output=True, text=True, check=True).stdout
    out = ['cuobjdump -sass %s  (sm_100a; %d lines)\n' % (os.path.relpath(LIB, ROOT), sass.count('\n')),
           'per kernel: mnemonic counts, then the first occurrence of each mnemonic\n']
    fn, counts, first = None, None, None

    def flush():
        if fn and counts and any(k.startswith(('UTC', 'UTMA', 'LDTM', 'UBLK', 'LDGMC')) for k in counts):
            out.append('\n== %s' % fn)
            out.append('   ' + '  '.join('%s x%d' % kv for kv in sorted(counts.items())))
            for k in sorted(first):
                out.append('      ' + first[k])
    for line in sass.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            flush()
            fn = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            fn = re.sub(r'\(.*', '', fn)
            counts, first = collections.Counter(), {}
            continue
        if counts is None:
            continue
        m = PAT.search(line)
        if m and '/*' in line:
            k = m.group(1)
            counts[k] += 1
            if k not in first:
                first[k] = re.sub(r'\s+', ' ', line.split('/* 0x')[0]).strip()
    flush()
    path = os.path.join(ROOT, 'profiles', 'r02_sass_excerpt.txt')
    open(path, 'w').write('\n'.join(out) + '\n')
    print(path, len(out), 'lines')


if __name__ == '__main__':
    sys.exit(main())
