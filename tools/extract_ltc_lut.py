"""Extract the reference's LTC lookup table (DATA, not code) into a binary fixture.

Reads /root/reference/include/ltc/ltc_isotropic.h:4,6,8 (three float[256] arrays = 8x8 RGBA32F,
uploaded as textures at include/viewer.hpp:321-327) and writes
optix_renderer_b200/data/ltc_isotropic_8x8.npy with shape (3, 8, 8, 4) float32.
Run in the build container only (the reference is not present on the GPU box).
"""
import os
import re
import sys

import numpy as np

ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
src = open(os.path.join(ref, "include/ltc/ltc_isotropic.h")).read()
tabs = []
for name in ("ltc_iso_1", "ltc_iso_2", "ltc_iso_3"):
    m = re.search(name + r"\[256\]\s*=\s*\{([^}]*)\}", src)
    vals = [float(t.strip().rstrip("f")) for t in m.group(1).split(",") if t.strip()]
    assert len(vals) == 256, (name, len(vals))
    tabs.append(np.array(vals, np.float32).reshape(8, 8, 4))
out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "optix_renderer_b200", "data",
                   "ltc_isotropic_8x8.npy")
np.save(out, np.stack(tabs))
print("wrote", out)
