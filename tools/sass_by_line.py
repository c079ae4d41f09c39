"""Static SASS instruction counts of one kernel by innermost source line, grouped under the kernel-level line that
inlines them. Input: `nvdisasm --print-line-info-inline x.cubin` text restricted to one function.
Usage: python tools/sass_by_line.py ext.sass <kernel-file-line, e.g. 303> [...]"""
import collections, re, sys

def main():
    path, tops = sys.argv[1], [int(a) for a in sys.argv[2:]]
    cnt = collections.Counter()
    group = []; last_was_file = False
    cur = top = None
    for ln in open(path):
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            ent = (m.group(1).split('/')[-1], int(m.group(2)))
            if not last_was_file: group = []
            group.append(ent); last_was_file = True
            cur, top = group[0], group[-1]
            continue
        last_was_file = False
        if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', ln):
            cnt[(top, cur)] += 1
    for t in tops:
        sel = {c: v for (tp, c), v in cnt.items() if tp[1] == t}
        print(f"--- kernel line {t}: {sum(sel.values())} instructions")
        for c in sorted(sel): print(f"  {c[0]}:{c[1]:<5} {sel[c]}")

main()
