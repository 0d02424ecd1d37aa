"""GPU debug aid: lanes where the EXACT kernels differ from the golden vectors."""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in ("mitsuba3-sunsky_b200", "tests"):
    sys.path.insert(0, os.path.join(REPO, p))
import numpy as np
import sunsky_b200 as ssk, common, emitters
for name in (sys.argv[1:] or list(common.CONFIGS)):
    g = emitters.golden(name); e = emitters.make(ssk, name, g=g)
    wav = g["wav"] if e.spectral else None
    out = e.eval(-g["wo"], wav); ref = g["eval"]
    bad = np.argwhere((out.view(np.uint32) != ref.view(np.uint32)) & ~((out == 0) & (ref == 0)))
    print(name, "eval mismatches", len(bad), "of", out.size)
    for c, i in bad[:8]:
        print("   ch", c, "lane", i, "got", out[c, i].hex() if hasattr(out[c,i],'hex') else out[c,i], float(out[c, i]), "want", float(ref[c, i]),
              "wav", None if wav is None else float(wav[c, i]), "wo", g["wo"][:, i])
    pdf = e.pdf_direction(g["wo"]); rp = g["pdf"]
    badp = np.nonzero((pdf.view(np.uint32) != rp.view(np.uint32)) & ~((pdf == 0) & (rp == 0)))[0]
    print(name, "pdf mismatches", len(badp))
    for i in badp[:5]:
        print("   lane", i, float(pdf[i]), float(rp[i]), g["wo"][:, i])
