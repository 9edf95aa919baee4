"""Development aid: a long differential run of the CUDA decoder against the CPU oracle.

    python tools/fuzz_campaign.py [first_seed] [count]

Prints one line per failing seed and a summary; exit status 1 if anything differed."""
import importlib
import random
import sys
import time

sys.path.insert(0, ".")
sys.path.insert(0, "oracle")
zk = importlib.import_module("node-zkstream_b200")
synth = importlib.import_module("node-zkstream_b200.synth")
import zkoracle  # noqa: E402

first = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
count = int(sys.argv[2]) if len(sys.argv) > 2 else 500
ctx = zk.Context(0)
bad = 0
frames = 0
t0 = time.time()
for seed in range(first, first + count):
    rng = random.Random(seed)
    kind = seed % 5
    if kind == 0:
        c = synth.fuzz_corpus(seed, nsessions=rng.randrange(1, 60), max_frames=rng.randrange(1, 120))
    elif kind == 1:
        c = synth.fuzz_corpus(seed, nsessions=rng.randrange(100, 400), max_frames=rng.randrange(20, 200),
                              p_mutate=rng.random() * 0.5, p_tail=rng.random() * 0.5, p_badlen=rng.random() * 0.3)
    elif kind == 2:
        c = synth.session_corpus(seed, nsessions=rng.randrange(1, 200), max_frames=rng.randrange(1, 150))
    elif kind == 4:
        # children vectors longer than the decode tile, with names built to mislead the element guess,
        # lying counts, and ordinary frames around them
        import struct
        c = synth.Corpus()
        parts, xm = [synth.frame(synth.connect_response(rng))], {}
        for i in range(rng.randrange(1, 6)):
            style = rng.randrange(4)
            n = rng.randrange(200, 3000)
            names = []
            for k in range(n):
                r = rng.random()
                if style == 0:
                    names.append(b"n%07d" % k + bytes(rng.randrange(0, 16)))
                elif style == 1:
                    names.append(bytes(rng.choice(b"abcdefghij0123456789") for _ in range(rng.randrange(0, 30))))
                elif style == 2:
                    names.append(bytes(rng.randrange(0, 40)) if r < 0.5 else b"\x00\x00\x00\x05abcde\x00\x00\x00\x03xyz")
                else:
                    names.append(bytes(rng.getrandbits(8) for _ in range(rng.choice([0, 1, 3, 5, 300, 700]) if r < 0.2 else rng.randrange(1, 20))))
            op = rng.choice([8, 12, 12])
            claimed = n + rng.choice([0, 0, 0, 1, -1, -n - 3, 5])
            body = synth.resp_header(i, 9 + i, rng.choice([0, 0, 0, -101])) + synth.be32(claimed)
            body += b"".join(struct.pack(">i", len(x) if x or rng.random() < 0.5 else -1) + x for x in names)
            if op == 12:
                body += synth.stat_bytes(rng, 0, n)
            parts.append(synth.frame(body))
            xm[i] = op
            if rng.random() < 0.5:
                parts.append(synth.frame(synth.resp_header(-2, 1, 0)))
        c.add(b"".join(parts), xidmap=xm)
    else:
        # payloads full of plausible frames: the splitter's guesses are wrong on purpose
        c = synth.Corpus()
        inner = synth.cfg2_mixed(1, rng.randrange(5, 60), seed=seed).chunks[1]
        parts, xm = [synth.frame(synth.connect_response(rng))], {}
        for i in range(rng.randrange(1, 40)):
            cut = rng.randrange(0, len(inner))
            data = inner[cut:cut + rng.randrange(0, 30000)]
            parts.append(synth.frame(synth.resp_header(i, i, 0) + synth.jbuf(data) + synth.stat_bytes(rng)))
            xm[i] = 4
        c.add(b"".join(parts), xidmap=xm)
        c.add(inner, xidmap={i: 4 for i in range(100)})
    d, off, fl, hs = c.arrays()
    xt = c.xidtable()
    g = ctx.decode_host(d, off, fl, hs, xt)
    o = zkoracle.decode(d, off, fl, hs, xt)
    frames += g.n_frames
    diff = g.diff(o)
    if diff:
        bad += 1
        print("seed %d kind %d: differs in %r" % (seed, kind, diff[:4]), flush=True)
print("%d corpora, %d frames, %d failures, %.0f s, %d splitter fix-up rounds" % (
    count, frames, bad, time.time() - t0, ctx.split_fixups))
sys.exit(1 if bad else 0)
