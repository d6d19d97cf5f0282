"""Second, independent restatement of the reference's DPF-PIR path in pure Python.

TEST INFRASTRUCTURE (see oracle/__init__.py).  It exists so that a misreading of the Go code in
oracle/pir_oracle.c does not go unnoticed: it was written from the Go source separately, uses a
different AES (the `cryptography` package, OpenSSL underneath) and Python big-int arithmetic with
explicit int64 wrapping.  tests/golden/make_golden.py runs both and commits the vectors only when
they agree byte-for-byte.  Slow: use on small domains only.

Reference lines followed: dpf/common.go:56-75, dpf/client.go:17-150, dpf/server.go:16-101,
db.go:75-174, slot.go:19-30, query.go:278-297.
"""
from cryptography.hazmat.primitives.ciphers import Cipher, algorithms, modes

MASK64 = (1 << 64) - 1


def _wrap(v):
    """Go int (int64) wrap-around."""
    v &= MASK64
    return v - (1 << 64) if v >> 63 else v


class Prf:
    """prf() of dpf/common.go:62-75 with three fixed-key AES-128 blocks (Matyas-Meyer-Oseas)."""

    def __init__(self, keys):
        self.enc = [Cipher(algorithms.AES(bytes(k)), modes.ECB()).encryptor() for k in keys[:3]]

    def __call__(self, x):
        x = bytes(x)
        out = bytearray()
        for e in self.enc:
            c = e.update(x)
            out += bytes(a ^ b for a, b in zip(c, x))
        return out  # 48 bytes


def go_varint(buf):
    """encoding/binary.Varint: Uvarint then zig-zag.  Returns 0 when no terminating byte is found."""
    x = 0
    s = 0
    ux = None
    for i, b in enumerate(buf):
        if i == 10:
            ux = 0
            break
        if b < 0x80:
            if i == 9 and b > 1:
                ux = 0
            else:
                ux = x | (b << s)
            break
        x |= (b & 0x7F) << s
        s += 7
    if ux is None:
        ux = 0
    v = ux >> 1
    if ux & 1:
        v = ~v
    return _wrap(v)


def get_bit(n, pos, N=64):
    return (n & (1 << (N - pos))) >> (N - pos)


def generate_two_server(tape, num_bits, a, b):
    """ClientInitialize + GenerateTwoServer; tape = the bytes crypto/rand would have returned."""
    tape = bytes(tape)
    keys = [tape[16 * i:16 * i + 16] for i in range(4)]
    prf = Prf(keys)
    r1 = tape[64:81]
    k0 = {"SInit": bytes(r1[:16]), "TInit": r1[16] % 2, "CW": []}
    k1 = {"SInit": bytes(tape[81:97]), "TInit": (r1[16] % 2) ^ 1, "CW": []}
    s0, s1 = bytearray(k0["SInit"]), bytearray(k1["SInit"])
    t0, t1 = k0["TInit"], k1["TInit"]
    N = 64
    for i in range(num_bits):
        p0, p1 = prf(s0), prf(s1)
        t0L, t0R = p0[16] % 2, p0[33] % 2
        t1L, t1R = p1[16] % 2, p1[33] % 2
        abit = get_bit(a, N - num_bits + i + 1, N)
        keep, lose = (0, 17) if abit == 0 else (17, 0)
        cw = bytearray(18)
        for j in range(16):
            cw[j] = p0[lose + j] ^ p1[lose + j]
        cw[16] = t0L ^ t1L ^ abit ^ 1
        cw[17] = t0R ^ t1R ^ abit
        for j in range(16):
            s0[j] = p0[keep + j] ^ ((t0 * cw[j]) & 0xFF)
            s1[j] = p1[keep + j] ^ ((t1 * cw[j]) & 0xFF)
        tkeep = cw[17] if keep == 17 else cw[16]
        t0 = (p0[keep + 16] % 2) ^ ((tkeep * t0) & 0xFF)
        t1 = (p1[keep + 16] % 2) ^ ((tkeep * t1) & 0xFF)
        k0["CW"].append(bytes(cw))
        k1["CW"].append(bytes(cw))
    f = _wrap(b - go_varint(s0[:8]) + go_varint(s1[:8]))
    if t1 == 1:
        f = _wrap(-f)
    k0["FinalCW"] = k1["FinalCW"] = f
    return keys, k0, k1


def evaluate_2p(prf, num_bits, server_num, key, x):
    """(*Dpf).Evaluate2P, dpf/server.go:55-101."""
    s = bytearray(key["SInit"])
    t = key["TInit"]
    N = 64
    for i in range(num_bits):
        xbit = get_bit(x, N - num_bits + i + 1, N)
        out = prf(s)
        cw = key["CW"][i]
        count = 0
        for j in range(34):
            if j == 17:
                count = 0
            elif j == 33:
                count = 17
            out[j] ^= (t * cw[count]) & 0xFF
            count += 1
        if xbit == 0:
            s = out[:16]
            t = out[16] % 2
        else:
            s = out[17:33]
            t = out[33] % 2
    v = _wrap(go_varint(s[:8]) + t * key["FinalCW"])
    return v if server_num == 0 else _wrap(-v)


def go_mod2_is_even(res):
    """bits[i] = (int(math.Abs(float64(res % 2))) == 0), db.go:142; Go's % truncates toward zero."""
    r = abs(res) % 2
    return r == 0


def expand_shared_query(keys, key2p, share_number, db_size, group_size, keywords=None):
    h = db_size // group_size
    num_bits = 32 if keywords is not None else h.bit_length()
    prf = Prf(keys)
    bits = []
    for i in range(h):
        k = keywords[i] if keywords is not None else i
        bits.append(go_mod2_is_even(evaluate_2p(prf, num_bits, share_number, key2p, k)))
    return bits


def answer_with_bits(slots, slot_bytes, group_size, bits):
    """db.go:75-107; slots is a list of bytes objects."""
    res = [bytearray(slot_bytes) for _ in range(group_size)]
    for row, bit in enumerate(bits):
        if not bit:
            continue
        for col in range(group_size):
            idx = row * group_size + col
            if idx >= len(slots):
                break
            src = slots[idx]
            for j in range(min(len(res[col]), len(src))):
                res[col][j] ^= src[j]
    return b"".join(bytes(r) for r in res)
