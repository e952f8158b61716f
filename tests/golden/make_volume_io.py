"""Regenerates tests/golden/volume_io.json with the reference's own quantize() (oracle/_ref). Run in the build container."""
import hashlib, json, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import oracle
from test_volume_io import _synth16
ref = oracle.Ref()
out = {"quantize": []}
for dims, seed in (((16, 12, 10), 1), ((9, 9, 9), 2), ((32, 5, 3), 3)):
    data = np.frombuffer(_synth16(*dims, seed), dtype=np.uint8)
    out["quantize"].append({"dims": list(dims), "seed": seed,
                            "nonlinear": hashlib.sha256(ref.quantize16(data, *dims, False).tobytes()).hexdigest(),
                            "linear": hashlib.sha256(ref.quantize16(data, *dims, True).tobytes()).hexdigest()})
json.dump(out, open(os.path.join(HERE, "volume_io.json"), "w"), indent=1)
print(out)
