import torch
n = 2_460_000_000 // 2
x = torch.empty(n, dtype=torch.bfloat16, device="cuda")
y = torch.empty(n, dtype=torch.bfloat16, device="cuda")
def t(f, reps=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
ms = t(lambda: x.zero_()); print("memset 2.46 GB: %.3f ms  %.2f TB/s write" % (ms, 2.46e9 / ms / 1e9))
ms = t(lambda: x.fill_(1.5)); print("fill   2.46 GB: %.3f ms  %.2f TB/s write" % (ms, 2.46e9 / ms / 1e9))
ms = t(lambda: y.copy_(x)); print("copy   2.46 GB: %.3f ms  %.2f TB/s read+write" % (ms, 2 * 2.46e9 / ms / 1e9))
ms = t(lambda: x.sum()); print("sum    2.46 GB: %.3f ms  %.2f TB/s read" % (ms, 2.46e9 / ms / 1e9))
z = torch.empty(n // 4, dtype=torch.bfloat16, device="cuda")
ms = t(lambda: torch.add(x[: n // 4], 1, out=z)); print("add 0.6->0.6 GB: %.3f ms %.2f TB/s" % (ms, 2 * 0.615e9 / ms / 1e9))
