"""C5 latency, one graph branch vs pipelined groups, and a check that stepping groups of
streams separately gives the same columns as stepping them together."""


def main():
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch, mfcc_b200
    from bench import stream_latency
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device=dev)
    g = torch.Generator().manual_seed(1)
    hops = [torch.randint(-3000, 3000, (4096, 170), dtype=torch.int16, generator=g).to(dev) for _ in range(6)]
    op.reset(4096)
    want = [op.step(h).clone() for h in hops]
    op.reset(4096)
    for h, w in zip(hops, want):
        out = torch.empty_like(w)
        for a in range(0, 4096, 1000):
            b = min(a + 1000, 4096)
            op.step(h[a:b], out=out[a:b], first_stream=a)
        assert torch.equal(out, w)
    print("grouped stepping == whole stepping")
    for groups, zc in ((1, False), (1, "out"), (1, True), (1, "out"), (1, False), (1, "out")):
        r = stream_latency(op, dev, 4096, 2000, groups=groups, zero_copy=zc)
        print(groups, zc, {k: round(v, 1) for k, v in r.items() if k.endswith("p50") or k.endswith("p99")})
    # zero-copy result equals the copied one
    hh = hops[0].cpu().pin_memory()
    oh = torch.empty((4096, 32), dtype=torch.int16).pin_memory()
    op.reset(4096)
    op.step(hh, out=oh)
    torch.cuda.synchronize()
    assert torch.equal(oh, want[0].cpu())
    print("zero-copy step == device step")


if __name__ == "__main__":
    main()
