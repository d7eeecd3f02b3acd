"""Device-side synthetic collections (same mutation model as tools/gen_synth.py / SURVEY.md 8(d)),
built with torch on the GPU so that the 5.9 Gbp workload is ready in seconds.  Bench tooling only.

reference = i.i.d. uniform ACGT; every copy independently: per position u~U(0,1):
u < d/3 substitute by another base, d/3 <= u < 2d/3 delete, 2d/3 <= u < d keep + insert one base.
The text is laid out as the reference's parser would produce it from a FASTA collection
(one leading empty document, then every copy followed by the 0x02 separator).
"""
import torch

ACGT = (65, 67, 71, 84)


def make_reference(n, gen, dev):
    lut = torch.tensor(ACGT, dtype=torch.uint8, device=dev)
    return lut[torch.randint(0, 4, (n,), generator=gen, device=dev)]


def mutate(ref, codes, div, gen, dev, keep=None):
    """keep: positions that are copied unchanged (separators of a batch of copies)"""
    n = ref.numel()
    lut = torch.tensor(ACGT, dtype=torch.uint8, device=dev)
    r = torch.rand(n, generator=gen, device=dev)
    if keep is not None:
        r[keep] = 1.0
    snp = r < div / 3
    dele = (r >= div / 3) & (r < 2 * div / 3)
    ins = (r >= 2 * div / 3) & (r < div)
    s = ref.clone()
    k = int(snp.sum())
    if k:
        s[snp] = lut[(codes[snp] + torch.randint(1, 4, (k,), generator=gen, device=dev)) % 4]
    reps = torch.ones(n, dtype=torch.int64, device=dev)
    reps[dele] = 0
    reps[ins] = 2
    t = torch.repeat_interleave(s, reps)
    ip = (torch.cumsum(reps, 0) - 1)[ins]
    if ip.numel():
        t[ip] = lut[torch.randint(0, 4, (ip.numel(),), generator=gen, device=dev)]
    return t


def apply_svs(s, gen, count, lo=1000, hi=50000):
    """count block events on a device string, the model of tools/gen_synth.apply_svs: length uniform
    in [lo, hi]; 30 % deletion, 30 % tandem duplication, 20 % reversal, 20 % translocation."""
    for _ in range(count):
        n = s.numel()
        u = torch.rand(3, generator=gen, device=s.device).cpu().tolist()
        L = lo + int(u[0] * (hi + 1 - lo))
        if L >= n // 2:
            L = max(1, n // 4)
        a = int(u[1] * (n - L))
        blk = s[a:a + L]
        if u[2] < 0.3:
            s = torch.cat([s[:a], s[a + L:]])
        elif u[2] < 0.6:
            s = torch.cat([s[:a + L], blk, s[a + L:]])
        elif u[2] < 0.8:
            s = torch.cat([s[:a], blk.flip(0), s[a + L:]])
        else:
            rest = torch.cat([s[:a], s[a + L:]])
            b = int(torch.rand(1, generator=gen, device=s.device).item() * (rest.numel() + 1))
            s = torch.cat([rest[:b], blk, rest[b:]])
    return s


def _codes(s):
    codes = torch.zeros(s.numel(), dtype=torch.int64, device=s.device)
    codes[s == 67] = 1
    codes[s == 71] = 2
    codes[s == 84] = 3
    return codes


def make_text(ref_bp, copies, div, seed, dev, pad=64, leading_empty=True, first_copy=0, ncopies=None, svs=0):
    """-> (ref uint8[ref_bp] on dev, text uint8[N + pad] on dev, N, ndocs).  Copies
    first_copy .. first_copy+ncopies-1 of the `copies`-copy collection (every copy has its own
    generator stream, so a shard of the collection is reproducible on any rank)."""
    g0 = torch.Generator(device=dev)
    g0.manual_seed(seed)
    ref = make_reference(ref_bp, g0, dev)
    codes = torch.zeros(ref_bp, dtype=torch.int64, device=dev)
    codes[ref == 67] = 1
    codes[ref == 71] = 2
    codes[ref == 84] = 3
    if ncopies is None:
        ncopies = copies - first_copy
    cap = int(ncopies * (ref_bp * (1 + div) + 64 + svs * 50000) + 64 + pad)
    text = torch.empty(cap, dtype=torch.uint8, device=dev)
    w = 0
    ndocs = 0
    if leading_empty and first_copy == 0:
        text[0] = 2
        w = 1
        ndocs = 1
    for k in range(first_copy, first_copy + ncopies):
        g = torch.Generator(device=dev)
        g.manual_seed(seed * 1000003 + k + 1)
        if svs:
            base = apply_svs(ref, g, svs)
            t = mutate(base, _codes(base), div, g, dev)
            del base
        else:
            t = mutate(ref, codes, div, g, dev)
        L = t.numel()
        if w + L + 1 + pad > cap:
            raise RuntimeError("text buffer too small")
        text[w:w + L] = t
        text[w + L] = 2
        w += L + 1
        ndocs += 1
        del t
    text[w:w + pad] = 2
    return ref, text, w, ndocs


def make_text_many(ref_bp, copies, div, seed, dev, pad=64, first_copy=0, ncopies=None, batch=2000):
    """Same layout as make_text for collections of very many short copies (config 4: 300,000 copies of
    a 30 kbp reference): copies are mutated `batch` at a time as one string whose separators are
    kept.  first_copy must be a multiple of `batch` (every batch has its own generator stream)."""
    assert first_copy % batch == 0
    g0 = torch.Generator(device=dev)
    g0.manual_seed(seed)
    ref = make_reference(ref_bp, g0, dev)
    if ncopies is None:
        ncopies = copies - first_copy
    unit = torch.cat([ref, torch.full((1,), 2, dtype=torch.uint8, device=dev)])
    ucodes = _codes(unit)
    cap = int(ncopies * (ref_bp * (1 + div) + 2) + 4096 * 64 + 64 + pad)
    text = torch.empty(cap, dtype=torch.uint8, device=dev)
    w = 0
    ndocs = 0
    if first_copy == 0:
        text[0] = 2
        w = 1
        ndocs = 1
    for b0 in range(first_copy, first_copy + ncopies, batch):
        k = min(batch, first_copy + ncopies - b0)
        g = torch.Generator(device=dev)
        g.manual_seed(seed * 1000003 + b0 // batch + 1)
        base = unit.repeat(k)
        keep = torch.arange(ref_bp, base.numel(), ref_bp + 1, device=dev)
        t = mutate(base, ucodes.repeat(k), div, g, dev, keep=keep)
        L = t.numel()
        if w + L + pad > cap:
            raise RuntimeError("text buffer too small")
        text[w:w + L] = t
        w += L
        ndocs += k
        del t, base
    text[w:w + pad] = 2
    return ref, text, w, ndocs
