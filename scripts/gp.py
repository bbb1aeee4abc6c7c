"""Diagnostic: total and the ops matching a pattern (default: gather convs) out of scripts/eval_profile.py's last line (stdin)."""
import re
import sys
pat = sys.argv[1] if len(sys.argv) > 1 else "_gather"
line = sys.stdin.read().strip().splitlines()[-1]
items = [(n, v, c) for n, v, c in re.findall(r'(\S+) (\d+\.\d+)x(\d+)', line) if re.search(pat, n)]
tot = re.search(r'total (\d+\.\d+)', line)
print(tot.group(1), ' '.join(f'{n.replace("conv_", "")}={v}x{c}' for n, v, c in items))
