"""Cases shared by tests/test_reform.py and tests/golden/make_reform_golden.py."""
from poregen_b200 import _capi as capi
from poregen_b200.synth import SynthConfig

# (label, generator config, reform arguments). The reference only accepts stride 5 (reform.cpp:22, 241).
CASES = [
    ("k9m0_paf", SynthConfig(n_reads=12, median_samples=3000, mode=capi.PGM_MODE_MOVES_I8, stride=5, seed=401), ["-k", "9", "-m", "0", "-c"]),
    ("k9m0_tsv", SynthConfig(n_reads=12, median_samples=3000, mode=capi.PGM_MODE_MOVES_I8, stride=5, seed=401), ["-k", "9", "-m", "0"]),
    ("k9m8_paf", SynthConfig(n_reads=12, median_samples=3000, mode=capi.PGM_MODE_MOVES_I8, stride=5, seed=402), ["-k", "9", "-m", "8", "-c"]),
    ("k5m2_tsv", SynthConfig(n_reads=20, median_samples=2000, mode=capi.PGM_MODE_MOVES_I8, stride=5, trim_max=40, seed=403), ["-k", "5", "-m", "2"]),
    ("k1m0_paf", SynthConfig(n_reads=20, median_samples=2000, mode=capi.PGM_MODE_MOVES_I8, stride=5, trim_max=40, seed=404), ["-k", "1", "-m", "0", "-c"]),
    ("k6m5_paf", SynthConfig(n_reads=30, median_samples=1500, mode=capi.PGM_MODE_MOVES_I8, stride=5, n_frac=0.02, seed=405),
     ["-k", "6", "-m", "5", "-c"]),
]
