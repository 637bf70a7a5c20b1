"""CPU oracle for the delay-spectrum Gibbs sampler.

TEST INFRASTRUCTURE ONLY.  Nothing under ``draco_b200/`` imports this package; it is
used by ``tests/``, by ``__graft_entry__.smoke()`` and by the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` as the checker / timed CPU baseline.

Parity status: PINNED.  ``tests/golden/*.npz`` hold outputs of the reference's own
functions (``/root/reference/draco/analysis/delay.py:1475-2201`` executed in the build
container by ``tests/golden/make_golden.py`` through ``oracle/ref_loader.py``); the
oracle is checked against those vectors in ``tests/test_oracle.py``.  The reference has
no tests or golden vectors of its own for this path.
"""

from .gibbs_oracle import (  # noqa: F401
    InjectedDraws,
    RecordingDraws,
    cosine_sum_window,
    cut_data,
    calculate_delays,
    dft_c2c,
    dft_c2r,
    dft_complex,
    dft_r2c,
    gibbs_chain,
    gibbs_estimate,
    invert_no_zero,
    whiten_inputs,
    wiener_filter,
)
from .cross_oracle import (  # noqa: F401,E402
    CrossInjectedDraws,
    CrossRecordingDraws,
    complex_wishart,
    cross_gibbs_chain,
    cross_whiten_inputs,
    standard_complex_normal,
    standard_complex_wishart,
)
from .transform_oracle import delay_spectrum_fft, delay_spectrum_to_power_spectrum  # noqa: F401,E402
