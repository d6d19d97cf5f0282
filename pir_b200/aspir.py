"""Mirror of the reference's secret-shared ASPIR audit (aspir.go:215-295, query.go:103-116): the second
caller of the answer path.  Two (or more) servers each answer the SAME DPF query over a database of
per-item auth keys (group size 1) and XOR their share of the client's auth token into the result; the
audit shares XOR to zero iff the client knew the key of the item it asked for.

Server side (`GenerateAuditForSharedQuery*`) runs on the GPU through pir_audit_shared_query*; the client
side (`NewAuthTokenSharesForKey`, `CheckAudit`) is host arithmetic like in the reference.
"""
import ctypes as C

import numpy as np

from . import _capi
from .pir import DBMetadata, NewEmptySlot, NewRandomSlot, Slot, XorSlots

StatisticalSecurityBytes = 16  # test_constants.go


class AuditTokenShare:
    """aspir.go:217-219"""

    def __init__(self, T):
        self.T = T


class AuthTokenShare:
    """aspir.go:222-224"""

    def __init__(self, T):
        self.T = T


class AuthenticatedQueryShare:
    """aspir.go:31-34: a QueryShare plus this server's share of the auth token."""

    def __init__(self, QueryShare, AuthToken):
        self.QueryShare = QueryShare
        self.AuthToken = AuthToken

    def __getattr__(self, name):  # Go struct embedding: query.GroupSize etc. resolve to the QueryShare
        return getattr(self.QueryShare, name)


def NewAuthTokenSharesForKey(authKey, numShares):
    """aspir.go:227-243: shares[1:] random, shares[0] = key XOR all the others."""
    numBytes = len(authKey.Data)
    shares = [None] * numShares
    accumulator = NewEmptySlot(numBytes)
    for i in range(1, numShares):
        share = NewRandomSlot(numBytes)
        XorSlots(accumulator, share)
        shares[i] = AuthTokenShare(share)
    XorSlots(accumulator, authKey)
    shares[0] = AuthTokenShare(accumulator)
    return shares


def NewAuthenticatedIndexQueryShares(dbmd, index, authKey, groupSize, numShares, _tape=None):
    """query.go:104-116 (a DBMetadata method in the reference)."""
    queryShares = dbmd.NewIndexQueryShares(index, groupSize, numShares, _tape=_tape)
    authTokenShares = NewAuthTokenSharesForKey(authKey, numShares)
    return [AuthenticatedQueryShare(q, t) for q, t in zip(queryShares, authTokenShares)]


DBMetadata.NewAuthenticatedIndexQueryShares = NewAuthenticatedIndexQueryShares


def GenerateAuditForSharedQuery(keyDB, query, nprocs=1):
    """aspir.go:246-257.  The key database is expanded with group size 1 whatever the query says; the
    reference then restores the old group size, so a query whose group size is not 1 ends in
    "Invalid challenge ciphertext result" (aspir.go:275-277) — same here (PirError)."""
    q = query.QueryShare.to_pod(num_bits=len(query.QueryShare.KeyTwoParty.CW))
    tok = bytes(query.AuthToken.T.Data)
    n = keyDB.SlotBytes
    out = (C.c_uint8 * n)()
    _capi.check(_capi.lib().pir_audit_shared_query(keyDB._handle(), C.byref(q), tok, len(tok), out, n))
    return AuditTokenShare(Slot(bytes(out)))


def GenerateAuditForSharedQueryWithExpandedBits(keyDB, query, bits, nprocs=1):
    """aspir.go:261-283: reuse the bits of one expansion (e.g. the data database's) for the key database."""
    b = np.ascontiguousarray(np.asarray(bits, dtype=np.uint8))
    tok = bytes(query.AuthToken.T.Data)
    n = keyDB.SlotBytes
    out = (C.c_uint8 * n)()
    _capi.check(_capi.lib().pir_audit_shared_query_with_bits(
        keyDB._handle(), query.QueryShare.GroupSize, b.ctypes.data_as(C.POINTER(C.c_uint8)), b.size, tok, len(tok), out, n))
    return AuditTokenShare(Slot(bytes(out)))


def CheckAudit(*auditTokens):
    """aspir.go:286-299: true iff the audit shares XOR to all-zero."""
    res = NewEmptySlot(len(auditTokens[0].T.Data))
    for tok in auditTokens:
        XorSlots(res, tok.T)
    return not any(res.Data)
