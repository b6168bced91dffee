import time, torch, sys
sys.path.insert(0, '.')
from coslam_b200 import ops
dev = torch.device('cuda:0')
n = (1 << 22) * 43
torch.manual_seed(0)
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); x = ops.device_uniform(n, dev); e1.record()
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print('device stream: %d floats, kernel %.2f ms, host part %.2f ms, wall %.2f ms' % (n, e0.elapsed_time(e1), (t1 - t0) * 1e3, (t2 - t0) * 1e3))
    del x
saved = ops.DEVICE_STREAM_MIN; ops.DEVICE_STREAM_MIN = 1 << 60
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter(); x = ops.host_uniform((1 << 22, 43), dev); torch.cuda.synchronize(); t2 = time.perf_counter()
    print('host loop + pinned upload: wall %.2f ms' % ((t2 - t0) * 1e3)); del x
