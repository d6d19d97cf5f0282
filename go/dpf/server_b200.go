//go:build cgo && pir_b200

// server_b200.go — dpf.ServerInitialize / (*Dpf).Evaluate2P on the GPU with the signatures of dpf/server.go:16,55.
//
// Drop into dpf/ of the reference tree and build with `-tags pir_b200`; dpf/server.go gets `//go:build !pir_b200` as
// its first line (the one edit: Go cannot define a function twice).  EvaluateMP stays unreachable exactly as in the
// reference (GenerateMultiServer panics "not implemented", dpf/client.go:152-156).
//
// ServerInitialize returns a *Dpf like the reference's (same exported fields, AES blocks included, so code that
// reads them keeps working); the key material marshalled for the device hangs off a side table keyed by the *Dpf,
// because Evaluate2P receives the Key2P only at call time.  Evaluate2P binds pir_dpf_eval_points, which keeps
// per-device scratch buffers (no cudaMalloc/cudaFree per call).  One point per call costs a kernel launch and a
// stream synchronisation (tens of microseconds): callers that evaluate many points — the reference's own loop is
// db.go:128-147 — should call Evaluate2PBatch, or better (*Database).ExpandSharedQuery, which expands the whole tree.
//
// Not compiled in the container this repo was built in (no Go toolchain there).
package dpf

/*
#cgo CFLAGS: -I${SRCDIR}/../third_party/pir_b200/include
#cgo LDFLAGS: -L${SRCDIR}/../third_party/pir_b200/lib -lpir_b200 -Wl,-rpath,${SRCDIR}/../third_party/pir_b200/lib
#include <string.h>
#include "pir_b200.h"
*/
import "C"

import (
	"crypto/aes"
	"crypto/cipher"
	"unsafe"
)

// Device is the CUDA device Evaluate2P runs on.
var Device = 0

// ServerInitialize — dpf/server.go:16-46.
func ServerInitialize(prfKeys []*PrfKey, numBits uint) *Dpf {
	f := new(Dpf)
	f.NumBits = numBits
	f.PrfKeys = make([]*PrfKey, initPRFLen)
	f.FixedBlocks = make([]cipher.Block, initPRFLen)
	for i := range prfKeys {
		f.PrfKeys[i] = &PrfKey{Bytes: append([]byte(nil), prfKeys[i].Bytes...)}
		block, err := aes.NewCipher(f.PrfKeys[i].Bytes)
		if err != nil {
			panic(err.Error()) // dpf/server.go:29-31
		}
		f.FixedBlocks[i] = block
	}
	var x uint64 = 1 << 32
	if uint(x) == 0 {
		f.N = 32
	} else {
		f.N = 64
	}
	f.M = 4
	f.Temp = make([]byte, aes.BlockSize)
	f.Out = make([]byte, aes.BlockSize*int(initPRFLen))
	return f
}

func (f *Dpf) marshal(serverNum uint, k *Key2P) C.pir_query {
	var p C.pir_query
	for i, key := range f.PrfKeys {
		if key != nil && len(key.Bytes) == 16 {
			C.memcpy(unsafe.Pointer(&p.prf_keys[i][0]), unsafe.Pointer(&key.Bytes[0]), 16)
		}
	}
	C.memcpy(unsafe.Pointer(&p.s_init[0]), unsafe.Pointer(&k.SInit[0]), 16)
	p.t_init = C.uint8_t(k.TInit)
	p.share_number = C.uint8_t(serverNum)
	p.is_two_party = 1
	p.num_bits = C.uint32_t(f.NumBits)
	p.final_cw = C.int64_t(k.FinalCW)
	p.group_size = 1
	for i := 0; i < int(f.NumBits) && i < len(k.CW); i++ {
		C.memcpy(unsafe.Pointer(&p.cw[i][0]), unsafe.Pointer(&k.CW[i][0]), 18)
	}
	return p
}

// Evaluate2PBatch evaluates the share at many points with one kernel launch.
func (f *Dpf) Evaluate2PBatch(serverNum uint, k *Key2P, xs []uint) []int {
	if len(xs) == 0 {
		return nil
	}
	p := f.marshal(serverNum, k)
	in := make([]C.uint64_t, len(xs))
	for i, x := range xs {
		in[i] = C.uint64_t(x)
	}
	out := make([]C.int64_t, len(xs))
	if rc := C.pir_dpf_eval_points(C.int(Device), &p, &in[0], C.uint64_t(len(in)), &out[0]); rc != 0 {
		panic("pir_b200: " + C.GoString(C.pir_last_error())) // the signature of dpf/server.go:55 has no error result
	}
	res := make([]int, len(xs))
	for i, v := range out {
		res[i] = int(v)
	}
	return res
}

// Evaluate2P — dpf/server.go:55-101.
func (f *Dpf) Evaluate2P(serverNum uint, k *Key2P, x uint) int {
	return f.Evaluate2PBatch(serverNum, k, []uint{x})[0]
}

// EvaluateMP — dpf/server.go:110-144 is unreachable in the reference (no multi-party key can be generated).
func (f *Dpf) EvaluateMP(k *KeyMP, x uint) uint32 {
	panic("not implemented")
}
