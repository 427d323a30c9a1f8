"""JSON codec: Jackson / Float.toString known answers for the oracle, and the product codec (libsib.so, CPU-side
functions only -- no GPU needed) against the oracle bit for bit.
Reference call sites: InferenceBolt.java:76-77 (decode), :89-91 (encode); InstObj.java:7-16, PredObj.java:8-17."""
import struct

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st


def f32(x):
    return np.float32(x)


def bits(x):
    return struct.unpack("<I", struct.pack("<f", float(x)))[0]


# text -> expected float32 bits, by Jackson's rule: Double.parseDouble then (float)
DECODE_KAT = [
    ("1", 0x3F800000), ("1.0", 0x3F800000), ("-0.0", 0x80000000), ("0", 0x00000000),
    ("0.1", 0x3DCCCCCD), ("1e-5", bits(np.float32(1e-5))), ("1.0E-5", bits(np.float32(1e-5))),
    ("1E+2", 0x42C80000), ("255", 0x437F0000), ("16777217", 0x4B800000),  # 2^24+1 ties to even
    ("3.4028235E38", 0x7F7FFFFF), ("1e39", 0x7F800000), ("1e-46", 0x00000000), ("1.4E-45", 0x00000001),
    ("123456789012345678901234567890", bits(np.float32(1.2345678901234568e29))),
    ("0.000000000000000000000000000000000000000000001401298464324817", 0x00000001),
    # double-rounding witness: just above the midpoint of two floats, but the excess is below double precision, so
    # text->double lands exactly on the midpoint and (float) ties to even: 1.0. A direct text->float parse gives
    # 1.0000001 (0x3F800001). Jackson goes through double, so 0x3F800000 is the reference answer.
    ("1.00000005960464477550", 0x3F800000),
]


@pytest.mark.parametrize("text,want", DECODE_KAT)
def test_decode_known_answers(oracle, sib, text, want):
    payload = ('{"instances":[[[[%s]]]]}' % text).encode()
    for dec in (oracle.json_decode, sib.native.json_decode):
        got = dec(payload)
        assert isinstance(got, np.ndarray), (text, got)
        assert got.shape == (1, 1, 1, 1)
        assert bits(got.reshape(-1)[0]) == want, (dec.__module__, text, hex(bits(got.reshape(-1)[0])), hex(want))


def test_double_rounding_witness_is_real():
    # python float() is strtod; np.float32(str) parses via double too. A direct float parse would differ:
    assert np.float32(float("1.00000005960464477550")) == np.float32(1.0)


# java.lang.Float.toString known outputs
TOSTRING_KAT = [
    (1.0, "1.0"), (0.001, "0.001"), (1.0e-4, "1.0E-4"), (9999999.0, "9999999.0"), (1.0e7, "1.0E7"),
    (0.1, "0.1"), (0.3, "0.3"), (100.0, "100.0"), (1.0e10, "1.0E10"), (123456.7, "123456.7"),
    (3.4028235e38, "3.4028235E38"), (1.17549435e-38, "1.1754944E-38"),  # JDK>=19 shortest form; JDK 8 printed the 9-digit 1.17549435E-38 (JDK-4511638)
    (1.4e-45, "1.4E-45"),
    (0.0, "0.0"), (-0.0, "-0.0"), (-1.5, "-1.5"), (1.23456789e8, "1.2345679E8"), (0.5, "0.5"),
    (2.0e-3, "0.002"), (9.999999e-4, "9.999999E-4"), (1.0e-3, "0.001"), (16777216.0, "1.6777216E7"),
    (0.33333334, "0.33333334"), (1.0e-10, "1.0E-10"), (float("nan"), '"NaN"'), (float("inf"), '"Infinity"'),
    (float("-inf"), '"-Infinity"'),
    # exact ties: the float lies midway between the two shortest decimals; Float.toString (JDK >= 19 specification),
    # Ryu and std::to_chars take the even last digit. The oracle once broke these ties by double rounding noise
    # (found by tests/fuzz_codec_diff.cpp: 2^-12 printed as 2.4414063E-4).
    (2.0 ** -12, "2.4414062E-4"), (0.00146484375, "0.0014648438"), (0.00732421875, "0.0073242188"),
    (0.244140625, "0.24414062"), (2.44140625, "2.4414062"), (24.4140625, "24.414062"), (0.00439453125, "0.0043945312"),
]


@pytest.mark.parametrize("x,want", TOSTRING_KAT)
def test_float_to_string_known_answers(oracle, sib, x, want):
    assert oracle.float_to_string(x) == want
    assert sib.native.float_to_string(x) == want


@settings(max_examples=3000, deadline=None)
@given(st.floats(width=32, allow_nan=False, allow_infinity=False))
def test_float_to_string_product_equals_oracle_and_round_trips(x):
    import oracle
    import sib_b200 as sib
    a = oracle.float_to_string(x)
    b = sib.native.float_to_string(x)
    assert a == b, (x, a, b)
    assert np.float32(float(a.replace("E", "e"))) == np.float32(x)
    assert ("E" in a) == (not (1e-3 <= abs(np.float32(x)) < 1e7) and x != 0)


@settings(max_examples=300, deadline=None)
@given(st.lists(st.floats(min_value=0.0, max_value=1.0, width=32), min_size=10, max_size=10))
def test_probability_rows_encode_identically(row):
    import oracle
    import sib_b200 as sib
    p = np.array([row, row[::-1]], np.float32)
    assert oracle.json_encode(p) == sib.native.json_encode(p)


def test_encode_schema(oracle, sib):
    p = np.array([[0.5, 0.25, 1e-8, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.25]], np.float32)
    want = b'{"predictions":[[0.5,0.25,1.0E-8,0.0,0.0,0.0,0.0,0.0,0.0,0.25]]}'
    assert oracle.json_encode(p) == want
    assert sib.native.json_encode(p) == want


def test_payload_round_trip(oracle, sib, presets):
    ms = sib.modelspec
    for name in ("mnist-lenet5", "cifar-lenet5"):
        m = presets[name]
        x = ms.synthetic_images(m, 7, 2)
        text = ms.instances_json(x)
        for dec in (oracle.json_decode, sib.native.json_decode):
            got = dec(text)
            assert got.shape == x.shape
            assert got.tobytes() == x.tobytes()
    assert 9000 < len(ms.instances_json(ms.synthetic_images(presets["mnist-lenet5"], 0, 1))) < 13000
    assert 33000 < len(ms.instances_json(ms.synthetic_images(presets["cifar-lenet5"], 0, 1))) < 42000


_num = st.one_of(
    st.floats(width=32, allow_nan=False, allow_infinity=False).map(lambda v: repr(float(v))),
    st.integers(-10**6, 10**6).map(str),
    st.decimals(allow_nan=False, allow_infinity=False, places=12, min_value=-1000, max_value=1000).map(str),
    st.floats(allow_nan=False, allow_infinity=False).map(lambda v: repr(v)),
)


@settings(max_examples=500, deadline=None)
@given(st.lists(_num, min_size=6, max_size=6), st.sampled_from(["", " ", "\n\t "]))
def test_random_numbers_decode_identically(nums, ws):
    import oracle
    import sib_b200 as sib
    body = "[[[%s,%s,%s],[%s,%s,%s]]]" % tuple(nums)
    text = ('%s{%s"instances"%s:%s[%s]%s}' % (ws, ws, ws, ws, body.replace(",", "," + ws), ws)).encode()
    a = oracle.json_decode(text)
    b = sib.native.json_decode(text)
    assert isinstance(a, np.ndarray) and isinstance(b, np.ndarray), (text, a, b)
    assert a.shape == b.shape == (1, 1, 2, 3)
    assert a.tobytes() == b.tobytes(), (text, a, b)
    want = np.array([np.float32(float(s)) for s in nums], np.float32)  # float(s) is strtod; then narrow
    assert a.tobytes() == want.tobytes()


EJSON, ESHAPE = -1, -2
ERROR_CASES = [
    (b"", EJSON), (b"{", EJSON), (b"[]", EJSON), (b'{"instances":[[[[1.0]]]]', EJSON),
    (b'{"instances":[[[[1.0,]]]]}', EJSON), (b'{"instances":[[[[01]]]]}', EJSON), (b'{"instances":[[[[.5]]]]}', EJSON),
    (b'{"instances":[[[[1.]]]]}', EJSON), (b'{"instances":[[[[+1]]]]}', EJSON), (b'{"instances":[[[[NaN]]]]}', EJSON),
    (b'{"instance":[[[[1.0]]]]}', EJSON),                      # unknown property -> UnrecognizedPropertyException
    (b'{"instances":[[[[1.0]]]],"extra":1}', EJSON),
    (b'{"instances":[[[1.0]]]}', EJSON),                       # rank 3: number where float[] expected
    (b'{"instances":[[[[[1.0]]]]]}', EJSON),                   # rank 5: array where float expected
    (b'{"instances":[[[["a"]]]]}', EJSON),
    (b'{"instances":{"a":1}}', EJSON),
    (b'{"instances":[[[[1.0,2.0],[3.0]]]]}', ESHAPE),          # ragged -> Tensor.create IllegalArgumentException
    (b'{"instances":[[[[1.0]],[[2.0],[3.0]]]]}', ESHAPE),
    (b'{"instances":null}', ESHAPE),                           # NPE in Tensor.create
    (b'{}', ESHAPE),
    (b'{"instances":[]}', ESHAPE),
    (b'{"instances":[[[[]]]]}', ESHAPE),
    (b'{"instances":[[[null]]]}', ESHAPE),
]


@pytest.mark.parametrize("text,want", ERROR_CASES)
def test_decode_error_classes(oracle, sib, text, want):
    assert oracle.json_decode(text, cap=64) == want
    assert sib.native.json_decode(text, cap=64) == want


def test_decode_leniencies_shared_with_jackson(oracle, sib):
    for dec in (oracle.json_decode, sib.native.json_decode):
        # trailing tokens ignored (FAIL_ON_TRAILING_TOKENS off); null element of a primitive array reads 0.0
        got = dec(b'{"instances":[[[[1.5,null]]]]} trailing')
        assert got.reshape(-1).tolist() == [1.5, 0.0]
        # duplicate key: last one wins
        got = dec(b'{"instances":[[[[1.0]]]],"instances":[[[[2.0],[3.0]]]]}')
        assert got.shape == (1, 1, 2, 1) and got.reshape(-1).tolist() == [2.0, 3.0]
        # N > 1 is legal in the schema (InstObj.java:8)
        got = dec(b'{"instances":[[[[1]]],[[[2]]],[[[3]]]]}')
        assert got.shape == (3, 1, 1, 1)


def test_decode_capacity(oracle, sib):
    text = b'{"instances":[[[[1,2,3,4]]]]}'
    assert oracle.json_decode(text, cap=3) == -3
    assert sib.native.json_decode(text, cap=3) == -3
