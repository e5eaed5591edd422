"""Prints the per-phase deltas of the MMA-issuer / epilogue timelines written by DEXGG_CHAIN_CLOCKS (tools/chain_clocks.py)."""
import sys
lines = open(sys.argv[1]).read().strip().split('\n')
tile = int(sys.argv[2]) if len(sys.argv) > 2 else 2
recs = []
i = 0
while i < len(lines):
    if lines[i].startswith('kind'):
        recs.append((int(lines[i].split()[1]), [int(x) for x in lines[i + 1].split()[1:]], [int(x) for x in lines[i + 2].split()[1:]]))
        i += 3
    else:
        i += 1
for k in (0, 2):
    _, mma, epi = [x for x in recs if x[0] == k][-1]
    print('kind', k, 'events mma', len(mma), 'epi', len(epi))
    for name, ev in (('mma', mma), ('epi', epi)):
        # find the tile period: total events / tiles unknown -> print a window
        n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
        st = int(sys.argv[4]) if len(sys.argv) > 4 else 0
        print(' ', name, 'abs', ev[st:st + 3], 'deltas', [ev[j + 1] - ev[j] for j in range(st, min(st + n, len(ev) - 1))])
