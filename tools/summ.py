import sys, json
keys = sys.argv[1:] or ["ms_kernel_ms", "pack_ms", "ms_ms", "sort_ms", "emit_ms", "gbps"]
for line in sys.stdin:
    line = line.strip()
    if line.startswith("{"):
        last = line
d = json.loads(last)
print({k: round(d[k], 2) for k in keys if k in d})
