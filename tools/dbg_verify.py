import random, sys
sys.path.insert(0, "tests")
import genome_weaver_align_b200 as gw
import oracle_lib as O
rnd = random.Random(5)
pairs = []
for _ in range(400):
    m = rnd.randint(1, 250)
    q = "".join(rnd.choice("ACGT") for _ in range(m))
    ref = list("".join(rnd.choice("ACGT") for _ in range(rnd.randint(0, 8))) + q + "".join(rnd.choice("ACGT") for _ in range(rnd.randint(0, 8))))
    for _e in range(rnd.randint(0, 6)):
        p = rnd.randrange(len(ref)); r = rnd.random()
        if r < 0.5: ref[p] = rnd.choice("ACGTN")
        elif r < 0.75: ref.insert(p, rnd.choice("ACGT"))
        else: del ref[p]
    pairs.append(("".join(ref), q))
import os
if os.environ.get('EXTRA'):
    pairs.append(('', 'ACGT')); pairs.append(('ACGTACGT', 'A'))
if os.environ.get('PRE'):
    gw.alignBlockDetailed([('GATCTA', 'GCTATA'), ('TATAATAATA', 'TAATA')], 31)
got = gw.alignBlockDetailed(pairs, 31)
bad = []
for i, ((r, q), g) in enumerate(zip(pairs, got)):
    w = O.align_block_detailed(r, q, 31)
    if (g is None) != (w is None) or (g and (g["cigar"], g["pos"], g["nm"]) != (w["cigar"], w["pos"], w["nm"])):
        bad.append(i); print("BAD", i, len(r), len(q), g, w)
print("nbad", len(bad))
for i in bad[:5]:
    g1 = gw.alignBlockDetailed([pairs[i]], 31)[0]
    print("alone", i, g1)
    g2 = gw.alignBlockDetailed([pairs[i]] * 3, 31)
    print("x3", [x["cigar"] for x in g2])
