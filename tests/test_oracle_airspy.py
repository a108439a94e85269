"""Pins oracle/sdr_oracle_airspy.c against the reference's own iqconverter_float.c
(third_party/libairspy/libairspy/src, compiled into oracle/_ref/libsdr_ref.so): the complex float
stream must be identical bit for bit, with the converter state carried across calls of ragged lengths."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol

needs_ref = pytest.mark.skipif(not os.path.exists(ol.REF_SO), reason="oracle/_ref not built")


class AirspyState(C.Structure):
    _fields_ = [("avg", C.c_float), ("even", C.c_float * 24), ("odd", C.c_float * 12)]


def oracle_airspy():
    o = C.CDLL(ol.build_oracle())
    o.sdr_oracle_airspy_init.argtypes = [C.POINTER(AirspyState)]
    o.sdr_oracle_airspy_process.argtypes = [C.POINTER(AirspyState), C.c_void_p, C.c_int, C.c_void_p]
    return o


def synth_adc(n_real, seed, dc=37):
    """12-bit ADC words: tone near fs/4 (the Airspy IF) + noise + a DC offset."""
    rng = np.random.default_rng(seed)
    i = np.arange(n_real)
    x = 2048 + dc + 600 * np.cos(2 * np.pi * (0.25 + 0.013) * i) + 40 * rng.standard_normal(n_real)
    return np.clip(np.round(x), 0, 4095).astype(np.uint16)


@needs_ref
@pytest.mark.parametrize("lens", [[4096] * 6, [4, 8, 44, 48, 52, 4096, 12, 65536], [262144]])
def test_airspy_restatement_is_bit_exact(lens):
    r = C.CDLL(ol.REF_SO)
    r.sdr_ref_airspy_create.restype = C.c_void_p
    r.sdr_ref_airspy_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    r.sdr_ref_airspy_destroy.argtypes = [C.c_void_p]
    o = oracle_airspy()
    h = r.sdr_ref_airspy_create()
    st = AirspyState()
    o.sdr_oracle_airspy_init(C.byref(st))
    raw = synth_adc(sum(lens), 7)
    pos = 0
    for n in lens:
        chunk = np.ascontiguousarray(raw[pos:pos + n])
        a, b = np.zeros(n, np.float32), np.zeros(n, np.float32)
        r.sdr_ref_airspy_process(h, chunk.ctypes.data, n, a.ctypes.data)
        o.sdr_oracle_airspy_process(C.byref(st), chunk.ctypes.data, n, b.ctypes.data)
        assert a.tobytes() == b.tobytes(), (pos, n, np.flatnonzero(a != b)[:5])
        pos += n
    r.sdr_ref_airspy_destroy(h)
    # the converter moves the fs/4 tone to baseband: most of the power sits near DC afterwards
    z = b[0::2] + 1j * b[1::2]
    assert np.abs(z).mean() > 0.1
