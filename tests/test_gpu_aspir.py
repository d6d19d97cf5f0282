"""Secret-shared ASPIR audit on the GPU, mirroring aspir_test.go:60-111
(TestSharedASPIRCompleteness / TestSharedASPIRSoundness) through the Go-API mirror, plus bit-exact
agreement of each audit share with the oracle's answer XOR token."""
import numpy as np
import pytest

import oracle
from pir_b200 import _capi, aspir, pir
from tests.helpers import to_capi_query  # noqa: F401

pytestmark = pytest.mark.gpu

TestDBSize = 1 << 10
NumQueries = 6


def _keydb(gpu_lib):
    db = pir.GenerateRandomDB(TestDBSize, aspir.StatisticalSecurityBytes)
    db.Upload()
    return db


def test_shared_aspir_completeness(gpu_lib):
    keydb = _keydb(gpu_lib)
    rng = np.random.default_rng(5)
    for _ in range(NumQueries):
        index = int(rng.integers(0, TestDBSize))
        authKey = keydb.Slots[index]
        shares = keydb.NewAuthenticatedIndexQueryShares(index, authKey, 1, 2)
        audits = [aspir.GenerateAuditForSharedQuery(keydb, s, 1) for s in shares]
        assert aspir.CheckAudit(*audits), "secret shared ASPIR proof failed"
    keydb.Close()


def test_shared_aspir_soundness(gpu_lib):
    keydb = _keydb(gpu_lib)
    rng = np.random.default_rng(6)
    for _ in range(NumQueries):
        index = int(rng.integers(1, TestDBSize))
        authKey = keydb.Slots[0]  # the key of another item
        shares = keydb.NewAuthenticatedIndexQueryShares(index, authKey, 1, 2)
        audits = [aspir.GenerateAuditForSharedQuery(keydb, s, 1) for s in shares]
        assert not aspir.CheckAudit(*audits), "ASPIR proof succeeded with a false auth key"
    keydb.Close()


def test_audit_share_is_answer_xor_token_bit_exact(gpu_lib):
    """aspir.go:267-282 restated with the oracle: audit = oracle answer (group size 1) XOR token share."""
    keydb = _keydb(gpu_lib)
    flat = np.frombuffer(b"".join(bytes(s.Data) for s in keydb.Slots), dtype=np.uint8).reshape(TestDBSize, -1)
    index = 321
    shares = keydb.NewAuthenticatedIndexQueryShares(index, keydb.Slots[index], 1, 2)
    for s in shares:
        got = bytes(aspir.GenerateAuditForSharedQuery(keydb, s, 1).T.Data)
        oq = oracle.query_from_bytes(s.QueryShare.to_pod(num_bits=len(s.QueryShare.KeyTwoParty.CW)).to_bytes())
        want = oracle.answer(oq, flat, TestDBSize, keydb.SlotBytes) ^ np.frombuffer(bytes(s.AuthToken.T.Data), dtype=np.uint8)
        assert got == want.tobytes()
        # one expansion can serve both databases (aspir.go:261-283)
        bits = keydb.ExpandSharedQuery(s.QueryShare, 1)
        again = bytes(aspir.GenerateAuditForSharedQueryWithExpandedBits(keydb, s, bits, 1).T.Data)
        assert again == got
    keydb.Close()


def test_audit_rejects_group_size_other_than_one(gpu_lib):
    """aspir.go:275-277 after db.go:79-89: the restored group size makes len(res.Shares) != 1."""
    keydb = _keydb(gpu_lib)
    shares = keydb.NewAuthenticatedIndexQueryShares(3, keydb.Slots[3], 2, 2)
    with pytest.raises(_capi.PirError, match="Invalid challenge ciphertext result"):
        aspir.GenerateAuditForSharedQuery(keydb, shares[0], 1)
    keydb.Close()
