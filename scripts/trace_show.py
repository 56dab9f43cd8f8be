"""Developer tool: print a window of the CTA-0 timeline written by scripts/trace_run.py."""
import re, sys
path = sys.argv[1]; lo = int(sys.argv[2]); hi = int(sys.argv[3])
sec = {}; cur = None
for line in open(path).read().split('\n'):
    m = re.match(r'^(mma|tma|relay|epi\w+) (\d+)$', line)
    if m: cur = m.group(1); continue
    if cur and line.strip() and '@' in line:
        sec[cur] = [(int(x.split('@')[0], 16), int(x.split('@')[1])) for x in line.split()]
for k, v in sec.items():
    print(k, ' '.join(f"{tag:03x}@{t}" for tag, t in v if lo < t < hi))
