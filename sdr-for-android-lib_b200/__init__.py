"""B200-native IQ hot path of SDR-for-Android-lib: FFT + peak/strength estimator + SSB demod.

The product is the C-ABI shared library ``libsdr_b200.so`` (``include/sdr_b200.h``) built from
``csrc/`` for sm_100a.  This package is the thin host-side binding used by the tests and
``bench.py``; it holds no DSP arithmetic and has no CPU fallback: importing :mod:`binding`
raises if the library has not been built.
"""
from .binding import (  # noqa: F401
    Config, FrameResult, Plan, Context, SdrbError, lib, lib_path, plan_config, synth_host,
    CU8, CS16, CF32, RU16, STAGE_FFT, STAGE_SSB, STAGE_SPECTRAL_PULSE, STAGE_AUDIO_PULSE, format_bytes, format_dtype,
    POWER_LINEAR_F32, POWER_DB_F32, POWER_DB_F16, FrameEstimate, ESTIMATE_DTYPE, NOALLOC_MAX_FRAMES, estimate_noalloc,
    RESULT_DTYPE, PULSE_DTYPE, PulseResult, AudioPulseConfig, audio_pulse_config_default,
)
from .sharding import shard_streams, gather_records, strongest_peak  # noqa: F401
