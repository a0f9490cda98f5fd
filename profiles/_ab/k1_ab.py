"""A/B timing of two builds of the library on the same box: python k1_ab.py <lib.so> [steps]."""
import ctypes as C, sys, json, torch
lib = C.CDLL(sys.argv[1]); steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
lib.crb_last_error.restype = C.c_char_p
vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
lib.crb_index_create.argtypes = [C.POINTER(vp), vp, vp, vp, i64, i32, i32, i64, i32]
lib.crb_row_inv_norms.argtypes = [vp, vp, i64, i32, i32, vp]
lib.crb_search.argtypes = [vp, vp, i32, i32, i32, vp, vp, vp, vp, vp, vp, C.c_size_t, i32, vp]
n, d, k = 10_000_000, 768, 10
g = torch.Generator(device="cuda"); g.manual_seed(1)
rows = torch.empty((n, d), dtype=torch.float16, device="cuda")
for lo in range(0, n, 1 << 20):
    hi = min(n, lo + (1 << 20)); rows[lo:hi] = torch.randn((hi - lo, d), generator=g, device="cuda").to(torch.float16)
inv = torch.empty(n, dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
assert lib.crb_row_inv_norms(rows.data_ptr(), inv.data_ptr(), n, d, 0, st) == 0
h = vp(); assert lib.crb_index_create(C.byref(h), rows.data_ptr(), inv.data_ptr(), None, n, d, 0, 0, 0) == 0, lib.crb_last_error()
qs = [torch.randn((1, d), device="cuda") for _ in range(4)]
s = torch.empty((1, k), device="cuda"); r = torch.empty((1, k), dtype=torch.int64, device="cuda"); v = torch.empty(1, dtype=torch.int32, device="cuda")
ws = torch.zeros(64 << 20, dtype=torch.uint8, device="cuda")
def step(i):
    rc = lib.crb_search(h, qs[i % 4].data_ptr(), 1, k, 0, None, s.data_ptr(), r.data_ptr(), v.data_ptr(), None, ws.data_ptr(), ws.numel(), 1, st)
    assert rc == 0, lib.crb_last_error()
for i in range(10): step(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(steps): step(i)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(json.dumps({"lib": sys.argv[1].split("/")[-1], "steps": steps, "ms": round(ms, 4), "gbs": round(n * d * 2 / ms / 1e6, 1)}))
