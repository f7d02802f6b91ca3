"""Run scripts/micro/seg_bw.cu: effective bandwidth of 32 / 64 / 128 / 512-byte row segments (273 MB tensor)."""
import ctypes as C, os, torch
lib = C.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "libsegbw.so"))
rows = 8 * 200 * 334 // 32 * 32
buf = torch.zeros(rows * 512, dtype=torch.uint8, device="cuda")
sink = torch.zeros(16, dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for write in (1, 0):
    for seg in (32, 64, 128, 512):
        for blocks in (148 * 2,):
            ts = []
            for it in range(8):
                flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                rc = lib.seg_run(seg, write, C.c_void_p(buf.data_ptr()), C.c_longlong(rows), C.c_void_p(sink.data_ptr()), blocks, C.c_void_p(st))
                e.record(); torch.cuda.synchronize()
                assert rc == 0, rc
                ts.append(s.elapsed_time(e))
            m = sorted(ts)[len(ts) // 2]
            print(f"{'write' if write else 'read '} seg={seg:3d} B  {m * 1e3:7.1f} us  {rows * 512 / m / 1e9:5.2f} TB/s")
