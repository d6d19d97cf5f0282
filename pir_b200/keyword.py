"""Mirror of keyword.go: the sqrt-layered search tree whose second layer is a PIR database answered by the
two-server (secret-shared) GPU path.  Same names, argument meaning and error behaviour as the reference;
`PrivateQuery` is `SecondLayer.PrivateSecretSharedQuery` (keyword.go:76-81), i.e. the fused CUDA kernel with
GroupSize = sqrt(N) slots per row.  The Paillier variant (`PrivateEncryptedQuery`, keyword.go:84-89) is out of
scope.
"""
import math

from . import pir

padding = "\x00"  # keyword.go:9


class PrivateSqrtST:
    """keyword.go:17-24"""

    def __init__(self):
        self.FirstLayer = []
        self.SecondLayer = None
        self.NumKeys = 0
        self.SlotBytes = 0
        self.Width = 0
        self.Height = 0

    def BuildForData(self, data, device=0):
        """keyword.go:34-73.  Returns None on success and an error string otherwise (Go returns `error`).
        Like the reference, "sorted" means non-increasing: a pair data[i] < data[i+1] is rejected."""
        n = len(data)
        if math.sqrt(float(n)) != math.floor(math.sqrt(float(n))):
            return "length of data is not a perfect square"
        raw = [_bytes(s) for s in data]
        for i in range(n - 1):
            if raw[i] < raw[i + 1]:
                return "data not sorted"
        sqrtDim = int(math.sqrt(float(n)))
        firstLayerBoundries = [data[i] for i in range(sqrtDim, n, sqrtDim)]
        firstLayerBoundries.append(data[n - 1])
        slotBytes = pir.GetRequiredSlotSize(firstLayerBoundries)
        db = pir.Database(device)
        db.BuildForDataWithSlotSize(data, pir.GetRequiredSlotSize(data))
        self.FirstLayer = firstLayerBoundries
        self.SecondLayer = db
        self.SlotBytes = slotBytes
        self.NumKeys = n
        self.Width = sqrtDim
        self.Height = sqrtDim
        return None

    def PrivateQuery(self, query, nprocs=1):
        """keyword.go:76-81"""
        return self.SecondLayer.PrivateSecretSharedQuery(query, nprocs)

    def PrivateEncryptedQuery(self, query, nprocs=1):
        raise NotImplementedError("Paillier (cPIR) queries are outside the hot path this library replaces")

    def GetSecondLayerMetadata(self):
        """keyword.go:92-97"""
        return pir.DBMetadata(self.SecondLayer.SlotBytes, self.SecondLayer.DBSize)


def NewPrivateSqrtST():
    return PrivateSqrtST()


def _bytes(s):
    return s.encode("latin-1") if isinstance(s, str) else bytes(s)


def PadToPowerOf2(data):
    """keyword.go:100-113"""
    nextPower = int(math.pow(2, math.ceil(math.log2(float(len(data))))))
    return [data[i] if i < len(data) else padding for i in range(nextPower)]


def PadToSqrt(data):
    """keyword.go:116-131"""
    nextSqrt = int(math.ceil(math.sqrt(float(len(data)))))
    nextSqrt = nextSqrt * nextSqrt
    return [data[i] if i < len(data) else padding for i in range(nextSqrt)]
