import os, time, threading
for f in ["/sys/fs/cgroup/cpu.max", "/sys/fs/cgroup/cpu.stat", "/proc/loadavg", "/sys/fs/cgroup/cpuset.cpus.effective"]:
    try: print(f, open(f).read().strip().replace("\n", " | "))
    except Exception as e: print(f, e)
print("nproc", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)))
def spin(dur, out):
    t_end = time.perf_counter() + dur; last = time.perf_counter(); mx = 0; big = 0
    while True:
        now = time.perf_counter()
        gap = now - last
        if gap > mx: mx = gap
        if gap > 0.002: big += 1
        last = now
        if now > t_end: break
    out.append((mx * 1e3, big))
for nth in (1, 4):
    out = []; ths = [threading.Thread(target=spin, args=(3.0, out)) for _ in range(nth)]
    [t.start() for t in ths]; [t.join() for t in ths]
    print(nth, "python threads: (max gap ms, gaps>2ms)", out)
try: print("cpu.stat after:", open("/sys/fs/cgroup/cpu.stat").read().strip().replace("\n", " | "))
except Exception as e: print(e)
