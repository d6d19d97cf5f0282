"""Host-side mirrors of keyword.go / slot.go / db.go helpers (no GPU): padding, the sqrt search tree's
build rules and error strings, slot comparison and the database-dimension helpers."""
import math

from pir_b200 import keyword, pir


def _descending(n):
    data = keyword.PadToSqrt([str(i) for i in range(n)])
    return sorted(data, key=lambda s: s.encode("latin-1"), reverse=True)


def test_pad_to_sqrt_and_power_of_two():  # keyword.go:100-131
    for n in (1, 2, 3, 4, 5, 99, 100, 101, 1023, 1024, 1025):
        data = [str(i) for i in range(n)]
        p = keyword.PadToSqrt(data)
        r = math.isqrt(len(p))
        assert r * r == len(p) and (r - 1) * (r - 1) < n <= len(p)
        assert p[:n] == data and all(x == "\x00" for x in p[n:])
        q = keyword.PadToPowerOf2(data)
        assert len(q) & (len(q) - 1) == 0 and len(q) // 2 < n <= len(q) or n == 1
        assert q[:n] == data and all(x == "\x00" for x in q[n:])


def test_sqrtst_build_rules():  # keyword.go:34-73
    st = keyword.NewPrivateSqrtST()
    assert st.BuildForData([str(i) for i in range(10)]) == "length of data is not a perfect square"
    assert st.BuildForData(["1", "2", "3", "4"]) == "data not sorted"      # ascending is rejected
    data = _descending(200)                                                  # 225 entries
    assert st.BuildForData(data) is None
    assert (st.NumKeys, st.Width, st.Height) == (225, 15, 15)
    assert st.FirstLayer == [data[i] for i in range(15, 225, 15)] + [data[224]]
    assert len(st.FirstLayer) == 15
    assert st.SlotBytes == max(len(s) for s in st.FirstLayer)
    md = st.GetSecondLayerMetadata()
    assert (md.SlotBytes, md.DBSize) == (3, 225)
    assert [bytes(s.Data) for s in st.SecondLayer.Slots[:2]] == [data[0].encode().ljust(3, b"\0"), data[1].encode().ljust(3, b"\0")]


def test_slot_helpers():  # slot.go:56-63, 137-186
    assert pir.NewSlotFromString("ab", 4).Data == bytearray(b"ab\0\0")
    assert pir.NewSlotFromString("abcdef", 4).Data == bytearray(b"abcdef")      # never truncated
    assert pir.Slot(b"ab\0\0").ToString() == "ab" and pir.Slot(b"\0\0").ToString() == "\x00"
    assert pir.Slot(b"ab").Compare(pir.Slot(b"ab")) == 0
    assert pir.Slot(b"ab").Compare(pir.Slot(b"ab\0")) == -1 and pir.Slot(b"b").Compare(pir.Slot(b"ab")) == 1
    assert pir.GetRequiredSlotSize(["a", "abc", ""]) == 3 and pir.GetRequiredSlotSize([]) == 0


def test_dimension_helpers():  # db.go:399-446
    md = pir.DBMetadata(3, 1024)
    assert md.IndexToCoordinates(77, 32, 32) == (2, 13)
    assert md.GetDimentionsForDatabase(32, 1) == (32, 32)
    assert md.GetDimentionsForDatabase(32, 3) == (30, 34)       # 1024 / 96 = 10 columns of 3, 1024 / 30 = 34 rows
    assert md.GetDimentionsForDatabase(2048, 1) == (1, 1024)    # width clamps to 1
    assert md.GetSqrtOfDBSize() == 33
    assert pir.GetOptimalDBDimentions(3, 1024) == (19, 55)
    assert pir.GetOptimalWeightedDBDimentions(3, 1024, 2) == (9, 110)
    assert pir.GetOptimalDBDimentions(0, 0) == (0, 1)
