"""cProfile of retriever.invoke() (host side of the worker's call) over a 1M x 384 table."""
import cProfile, pstats, sys, io
sys.argv = [sys.argv[0]]
src = open("profiles/invoke_latency.py").read().split("for name, build in")[0]
exec(src)
r = f.for_chunk()
for i in range(10): r.invoke(str(1000 + 37 * i), filter={"namespace": "default"})
pr = cProfile.Profile(); pr.enable()
for i in range(100): r.invoke(str(2000 + 37 * i), filter={"namespace": "default"})
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(32); print(s.getvalue()[:6000])
