"""HBM read / write / copy bandwidth with torch's own kernels (context for the roofline numbers)."""
import torch


def timeit(f, n=20):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e-3


def main():
    nbytes = 1 << 30
    x = torch.empty(nbytes // 4, dtype=torch.float32, device="cuda").normal_()
    y = torch.empty_like(x)
    t = timeit(lambda: y.copy_(x))
    print(f"copy   : {2 * nbytes / t / 1e9:8.1f} GB/s (read + write)")
    t = timeit(lambda: y.zero_())
    print(f"memset : {nbytes / t / 1e9:8.1f} GB/s (write only)")
    t = timeit(lambda: y.fill_(1.5))
    print(f"fill   : {nbytes / t / 1e9:8.1f} GB/s (write only)")
    t = timeit(lambda: x.sum())
    print(f"sum    : {nbytes / t / 1e9:8.1f} GB/s (read only)")
    z = torch.empty(nbytes // 2, dtype=torch.float32, device="cuda")
    t = timeit(lambda: torch.add(x[: nbytes // 8], x[nbytes // 8 :], out=None) if False else z.copy_(x[: nbytes // 8].repeat(0) if False else z))
    a, b = x[: nbytes // 8], x[nbytes // 8 : nbytes // 4]
    o = torch.empty_like(a)
    t = timeit(lambda: torch.add(a, b, out=o))
    print(f"add    : {3 * a.numel() * 4 / t / 1e9:8.1f} GB/s (2 reads + 1 write, 128 MB operands)")


if __name__ == "__main__":
    main()
