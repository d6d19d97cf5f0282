"""keyword.go's sqrt search tree over the GPU answer path, mirroring TestKeywordQuerySqrtST
(keyword_test.go:25-102): a second-layer PIR query with GroupSize = sqrt(N) returns the whole row of
boundaries and the client finds the keyword's index in it.  Every share is also compared with the oracle."""
import numpy as np
import pytest

import oracle
from pir_b200 import keyword, pir

pytestmark = pytest.mark.gpu

NumTrials = 3


def _flat(db):
    return np.frombuffer(b"".join(bytes(s.Data) for s in db.Slots), dtype=np.uint8).reshape(db.DBSize, db.SlotBytes)


def test_keyword_query_sqrt_st(gpu_lib):
    rng = np.random.default_rng(11)
    for trial in range(NumTrials):
        numStrings = int(rng.integers(0, 1 << 10)) + 100
        data = keyword.PadToSqrt([str(i) for i in range(numStrings)])
        data = sorted(data, key=lambda s: s.encode("latin-1"), reverse=True)   # sort.Strings + argsort.ReverseStrings
        sqst = keyword.NewPrivateSqrtST()
        assert sqst.BuildForData(data) is None
        assert int(np.ceil(np.sqrt(len(data)))) == len(sqst.FirstLayer)
        flat = _flat(sqst.SecondLayer)
        picks = sorted({0, 1, len(data) - 1, len(data) - 2, *[int(x) for x in rng.integers(0, len(data), size=40)]})
        for i in picks:
            query = pir.NewSlotFromString(data[i], sqst.SlotBytes)
            rowIndex = 0
            for rowIndex, boundry in enumerate(sqst.FirstLayer):
                if data[i].encode("latin-1") > boundry.encode("latin-1"):
                    break
            shares = sqst.SecondLayer.NewIndexQueryShares(rowIndex, sqst.Height, 2)
            resA = sqst.PrivateQuery(shares[0], 4)
            resB = sqst.PrivateQuery(shares[1], 4)
            res = pir.Recover([resA, resB])
            assert len(res) == len(sqst.FirstLayer)
            colIndex = 0
            for colIndex, slot in enumerate(res):
                if slot.Compare(query) <= 0:
                    break
            index = rowIndex * sqst.Width + colIndex
            assert index == i or data[index] == data[i], (trial, i, index)
            if i == picks[0]:   # share-level parity with the oracle, once per trial
                for sh, got in zip(shares, (resA, resB)):
                    oq = oracle.query_from_bytes(sh.to_pod(num_bits=len(sh.KeyTwoParty.CW)).to_bytes())
                    want = oracle.answer(oq, flat, sqst.SecondLayer.DBSize, sqst.SecondLayer.SlotBytes)
                    assert b"".join(bytes(s.Data) for s in got.Shares) == want.tobytes()
        sqst.SecondLayer.Close()
