import cProfile, pstats, sys, io
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests/golden")
from draco_b200 import delayopt
from make_golden_nrml import make_nrml_case
N, M, T, maxiter = 2048, 1024, 64, 8
data, Ni, fsel = make_nrml_case(N, M, T, 20, 3)
delayopt.delay_power_spectrum_maxpost(data, N, Ni, None, window="nuttall", fsel=fsel, maxiter=2)
pr = cProfile.Profile(); pr.enable()
delayopt.delay_power_spectrum_maxpost(data, N, Ni, None, window="nuttall", fsel=fsel, maxiter=maxiter)
pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28); print(s.getvalue())
