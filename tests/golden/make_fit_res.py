"""Golden vectors for geometry.fit_resolution, produced by IMPORTING the reference's own
`fit_res` (src/selkies/legacy/resize.py:15-29) in the build container.  /root/reference does not
exist on the GPU box, so only the committed JSON travels.

    python tests/golden/make_fit_res.py
"""
import importlib.util
import json
import os
import random

REF = "/root/reference/src/selkies/legacy/resize.py"
spec = importlib.util.spec_from_file_location("ref_resize", REF)
mod = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mod)

cases = [(1920, 1080, 7680, 4320), (7680, 4320, 7680, 4320), (7681, 4321, 7680, 4320), (8192, 4096, 7680, 4320),
         (10000, 3000, 7680, 4320), (3000, 10000, 7680, 4320), (2560, 1600, 2560, 1600), (3840, 2160, 2560, 1600),
         (2561, 1440, 2560, 1600), (1366, 768, 7680, 4320), (15360, 8640, 7680, 4320), (7679, 4319, 7680, 4320),
         (7680, 100, 7680, 4320), (100, 4320, 7680, 4320)]
rng = random.Random(7)
for _ in range(40):
    cases.append((rng.randrange(2, 20000), rng.randrange(2, 12000), *rng.choice([(7680, 4320), (2560, 1600)])))
out = [{"in": list(c), "out": list(mod.fit_res(*c))} for c in cases]
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fit_res.json")
json.dump({"_source": "reference fit_res imported from " + REF, "vectors": out}, open(path, "w"), indent=0)
print("wrote", path, len(out))
