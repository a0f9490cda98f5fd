"""A few retriever.invoke() calls over the 1M x 384 chunk table, for an ncu launch list (kernel durations per traversal step)."""
import sys
sys.argv = [sys.argv[0]]
src = open("profiles/invoke_latency.py").read().split("for name, build in")[0]
exec(src)
r = f.for_chunk()
for i in range(6):
    r.invoke(str(2000 + 37 * i), filter={"namespace": "default"})
