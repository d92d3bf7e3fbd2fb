"""Tensor-list shapes of the BASELINE.json configs (SURVEY.md Appendix C shape table)."""


def gpt2_shapes(n_embd=768, n_layer=12, vocab=50257, n_pos=1024):
    """Parameter sizes of GPT2LMHeadModel(GPT2Config(n_embd, n_layer)) with the tied LM head."""
    e = n_embd
    sizes = [vocab * e, n_pos * e]
    for _ in range(n_layer):
        sizes += [e, e, e * 3 * e, 3 * e, e * e, e, e, e, e * 4 * e, 4 * e, 4 * e * e, e]
    sizes += [e, e]
    return sizes


GPT2_SMALL = gpt2_shapes()
GPT2_MEDIUM = gpt2_shapes(1024, 24)
assert len(GPT2_SMALL) == 148 and sum(GPT2_SMALL) == 124_439_808
assert len(GPT2_MEDIUM) == 292 and sum(GPT2_MEDIUM) == 354_823_168

CONVNET = [16 * 25, 16, 16, 16, 32 * 16 * 25, 32, 32, 32, 10 * 1568, 10]
assert sum(CONVNET) == 29_034


def many_small(total=1_300_000_000, lo=256, hi=262_144, seed=0):
    """BASELINE config 5: sizes log-uniform in [lo, hi], some not divisible by 8."""
    import math
    import random
    rng = random.Random(seed)
    sizes, acc = [], 0
    while acc < total:
        n = int(math.exp(rng.uniform(math.log(lo), math.log(hi))))
        sizes.append(n)
        acc += n
    return sizes
