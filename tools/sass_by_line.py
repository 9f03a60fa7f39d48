"""Static SASS instruction count per source line for one kernel.
cuobjdump -xelf all file.o; nvdisasm -g x.cubin > all.dis; python tools/sass_by_line.py all.dis <kernel substring> [top]"""
import re, sys
from collections import Counter
dis, key = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
inside = False
cur = None
cnt = Counter()
total = 0
for l in open(dis):
    if l.startswith(".text."):
        inside = key in l
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l):
        cnt[cur] += 1
        total += 1
print("total", total)
for (f, n), c in cnt.most_common(top):
    print("%6d  %s:%d" % (c, f, n))
