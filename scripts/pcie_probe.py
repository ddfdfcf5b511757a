"""PCIe floor for the end-to-end SpMV line: 64 MB up, 64 MB down, alone and at the same time."""
import time
import torch

n = 8_000_000
h_in = torch.empty(n, dtype=torch.float64).pin_memory()
h_out = torch.empty(n, dtype=torch.float64).pin_memory()
d_in = torch.empty(n, dtype=torch.float64, device="cuda")
d_out = torch.empty(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


def up():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def down():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    up()
    down()


def both_sync():
    both()
    torch.cuda.synchronize()


print("h2d 64MB ms", timed(up), "d2h 64MB ms", timed(down), "both back-to-back ms", timed(both), "both, synchronised each step ms", timed(both_sync))


# chunked: 8 pieces each way; (a) independent, (b) download i gated on upload i by an event
pieces = 8
step = n // pieces


def chunked(gated):
    evs = []
    with torch.cuda.stream(s1):
        for i in range(pieces):
            d_in[i * step:(i + 1) * step].copy_(h_in[i * step:(i + 1) * step], non_blocking=True)
            e = torch.cuda.Event()
            e.record(s1)
            evs.append(e)
    with torch.cuda.stream(s2):
        for i in range(pieces):
            if gated:
                s2.wait_event(evs[i])
            h_out[i * step:(i + 1) * step].copy_(d_out[i * step:(i + 1) * step], non_blocking=True)


def interleaved(gated):
    # issue order upload 0, download 0, upload 1, download 1, ... (the library's order)
    for i in range(pieces):
        with torch.cuda.stream(s1):
            d_in[i * step:(i + 1) * step].copy_(h_in[i * step:(i + 1) * step], non_blocking=True)
            e = torch.cuda.Event()
            e.record(s1)
        with torch.cuda.stream(s2):
            if gated:
                s2.wait_event(e)
            h_out[i * step:(i + 1) * step].copy_(d_out[i * step:(i + 1) * step], non_blocking=True)


print("chunked independent ms", timed(lambda: chunked(False)), "chunked gated ms", timed(lambda: chunked(True)),
      "interleaved independent ms", timed(lambda: interleaved(False)), "interleaved gated ms", timed(lambda: interleaved(True)))


def relayed(relay):
    # the library's chain: upload i (s1) -> event -> relay stream (waits, records) -> event -> download i (s2)
    for i in range(pieces):
        with torch.cuda.stream(s1):
            d_in[i * step:(i + 1) * step].copy_(h_in[i * step:(i + 1) * step], non_blocking=True)
            e = torch.cuda.Event()
            e.record(s1)
        relay.wait_event(e)
        e2 = torch.cuda.Event()
        e2.record(relay)
        with torch.cuda.stream(s2):
            s2.wait_event(e2)
            h_out[i * step:(i + 1) * step].copy_(d_out[i * step:(i + 1) * step], non_blocking=True)


s3 = torch.cuda.Stream()
print("relay through the legacy default stream ms", timed(lambda: relayed(torch.cuda.default_stream())),
      "relay through a third stream ms", timed(lambda: relayed(s3)))


# does a kernel that saturates HBM slow the copy engines down?
big_a = torch.empty(256 * 1024 * 1024 // 8, dtype=torch.float64, device="cuda")
big_b = torch.empty_like(big_a)


def both_under_load():
    with torch.cuda.stream(s3):
        for _ in range(12):          # ~12 x 80 us of device-to-device copy at full HBM rate
            big_b.copy_(big_a, non_blocking=True)
    both()


print("both directions while a copy kernel saturates HBM ms", timed(both_under_load), "(the 12 device copies alone ms", timed(lambda: [big_b.copy_(big_a) for _ in range(12)]), ")")
