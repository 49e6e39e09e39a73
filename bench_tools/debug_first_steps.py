import time, torch, sys
sys.path.insert(0, "/root/repo")
from dps_b200 import Config
from dps_b200.api import Job
for mode in ("async", "sync"):
    cfg = Config(mode=mode, workers=1, batch_size=128)
    job = Job(cfg, rank=0, world=1, device_index=0)
    x = torch.randn(128, 3, 32, 32).pin_memory(); y = torch.randint(0, 100, (128,)).pin_memory()
    ts = []
    for i in range(6):
        torch.cuda.synchronize(); t = time.time()
        job.worker.train_step(x, y).item()
        torch.cuda.synchronize(); ts.append(round((time.time() - t) * 1e3, 2))
    print(mode, "ms per step:", ts, "err", job.ps.error_bits(), job.ps.async_stats() if mode == "async" else "")
    job.close()
