import sys, os, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import native
L = native.lib()
L.b200_stream_probe.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p]
buf = torch.empty(2 << 30, dtype=torch.uint8, device="cuda")
sink = torch.zeros(1, dtype=torch.int64, device="cuda")
def run(total, stage, stages, grid, hint=1):
    st = torch.cuda.current_stream().cuda_stream
    nrot = buf.numel() // total
    for i in range(2):
        native.check(L.b200_stream_probe(buf.data_ptr(), total, stage, stages, grid, hint, sink.data_ptr(), st))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    reps = 10
    e0.record()
    for i in range(reps):
        native.check(L.b200_stream_probe(buf.data_ptr() + (i % nrot) * total, total, stage, stages, grid, hint, sink.data_ptr(), st))
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    print(f"total={total>>20}MB stage={stage>>10}KB x{stages} grid={grid} hint={hint}: {us:.1f} us {total/us/1e3:.0f} GB/s ({total/us/1e3/grid:.1f} GB/s/CTA)", flush=True)
big = 512 << 20
for stage, stages in ((16384, 6), (16384, 12), (32768, 3), (32768, 6), (8192, 12), (8192, 24), (65536, 3)):
    run(big, stage, stages, 148)
for grid in (1, 8, 32, 74, 148):
    run(64 << 20 if grid < 32 else big, 16384, 6, grid)
    run(64 << 20 if grid < 32 else big, 16384, 12, grid)
run(big, 16384, 6, 148, hint=0)
run(big, 16384, 6, 296)
run(big, 16384, 3, 296)
