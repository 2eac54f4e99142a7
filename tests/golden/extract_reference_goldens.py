"""Copies the golden force trajectories out of the reference's own test file into a JSON fixture.

Run in the build container (where /root/reference exists):
    python tests/golden/extract_reference_goldens.py
Source: /root/reference/tests/shard1/test_ivp.py (the np.array literals at lines 28-44, 65-89,
107-124, 150-168).  The fixture travels to the GPU box; the reference tree does not.
"""
import json
import re
from pathlib import Path

SRC = Path("/root/reference/tests/shard1/test_ivp.py")
OUT = Path(__file__).with_name("ivp_golden.json")

text = SRC.read_text()
arrays = [
    [float(v) for v in re.findall(r"-?\d+\.\d*(?:e-?\d+)?", body)]
    for body in re.findall(r"np\.array\(\s*\[(.*?)\]\)", text, flags=re.S)
]
# order of appearance in test_ivp.py
names = [
    "ding2003_with_fatigue", "ding2003",
    "ding2007_with_fatigue", "ding2007",
    "hmed2018_with_fatigue", "hmed2018",
    "pulse_mode_single", "pulse_mode_doublet_150_180", "pulse_mode_triplet_350_380",
]
assert len(arrays) == len(names), (len(arrays), len(names))
fixture = {
    "source": "reference tests/shard1/test_ivp.py (stim [0,0.1,0.2], sum_stim_truncation=3, final_time=0.3, RK4(10))",
    "decimal": 7,
    "pulse_width": [0.0003, 0.0004, 0.0005],
    "pulse_intensity": [50, 60, 70],
    "F": dict(zip(names, arrays)),
}
assert all(len(fixture["F"][k]) == 31 for k in names[:7]) and len(arrays[7]) == 30 and len(arrays[8]) == 30
OUT.write_text(json.dumps(fixture, indent=1))
print("wrote", OUT, {k: len(v) for k, v in fixture["F"].items()})
