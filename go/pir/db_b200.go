//go:build cgo && pir_b200

// db_b200.go — the B200 answer path behind the reference's own methods.
//
// Drop this file into the root of the reference tree (package pir), build with `-tags pir_b200`, and
// (*Database).PrivateSecretSharedQuery / ExpandSharedQuery / PrivateSecretSharedQueryWithExpandedBits run on the
// GPU with the signatures of db.go:68, 75 and 110.  Their callers — aspir.go:253,267 (GenerateAuditForSharedQuery),
// keyword.go:80 (PrivateSqrtST.PrivateQuery), db_test.go:32-37 — compile and run unchanged.
//
// The ONE edit to existing files: Go has no method override, so the three CPU methods (db.go:62-174) move, verbatim,
// into a file `db_shared_cpu.go` that starts with `//go:build !pir_b200` (INTEGRATION.md shows the two commands).
//
// State: the device copy of db.Slots hangs off an unexported side table keyed by the *Database (no field is added to
// the reference's struct).  It is created on the first query — flattening []*Slot into one slotBytes-pitched buffer,
// cgo cannot pass memory that holds Go pointers — and freed by a finalizer or db.ReleaseGPU().  The reference has no
// update API; code that mutates db.Slots after the first query calls db.InvalidateGPU().
//
// Errors: the reference's hot path never fails (db.go:106 returns nil).  A CUDA failure is returned as `error` where
// the signature has one.  ExpandSharedQuery (db.go:110) returns only []bool: a failure there is kept on the handle,
// the method returns nil, and the next call that has an error result reports it.
//
// Not compiled in the container this repo was built in (no Go toolchain there, none on the GPU box): the same C ABI
// is exercised from C99 (tests/cabi_smoke.c) and ctypes (pir_b200/_capi.py).
package pir

/*
#cgo CFLAGS: -I${SRCDIR}/third_party/pir_b200/include
#cgo LDFLAGS: -L${SRCDIR}/third_party/pir_b200/lib -lpir_b200 -Wl,-rpath,${SRCDIR}/third_party/pir_b200/lib
#include <stdlib.h>
#include <string.h>
#include "pir_b200.h"
*/
import "C"

import (
	"errors"
	"runtime"
	"sync"
	"unsafe"

	"github.com/sachaservan/pir/dpf"
)

// GPUDevices lists the CUDA devices a Database is row-sharded over (SURVEY 8e: GPU g holds rows
// [g*H/G, (g+1)*H/G)).  Set before the first query; default: device 0 only.
var GPUDevices = []int{0}

// GPUShardAlign is the largest GroupSize the databases of this process are queried with: shard boundaries are
// multiples of it, so that every query's rows fall into exactly one shard.
var GPUShardAlign = 1

type gpuDB struct {
	mu      sync.RWMutex // lifetime of srv only: queries hold it for reading, ReleaseGPU for writing.  The C library is
	srv     *C.pir_server //   thread-safe and keeps up to 8 single queries of different goroutines in flight per handle.
	errMu   sync.Mutex
	lastErr error // failure of a call whose signature has no error result (ExpandSharedQuery)
}

// takeErr returns and clears the error a call without an error result has left behind.
func (g *gpuDB) takeErr() error {
	g.errMu.Lock()
	defer g.errMu.Unlock()
	err := g.lastErr
	g.lastErr = nil
	return err
}

func (g *gpuDB) keepErr(err error) {
	g.errMu.Lock()
	g.lastErr = err
	g.errMu.Unlock()
}

var errReleased = errors.New("pir_b200: the device copy of this Database was released")

// keyed by the Database's ADDRESS (a uintptr does not keep the Database alive, so its finalizer can run; the finalizer
// removes the entry before the address can be handed out again)
var gpuTable sync.Map // uintptr(unsafe.Pointer(*Database)) -> *gpuDB

func cErr(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return errors.New("pir_b200: " + C.GoString(C.pir_last_error()))
}

// gpu returns the device copy of db, uploading it on first use.
func (db *Database) gpu() (*gpuDB, error) {
	if v, ok := gpuTable.Load(uintptr(unsafe.Pointer(db))); ok {
		return v.(*gpuDB), nil
	}
	g := &gpuDB{}
	devs := make([]C.int, len(GPUDevices))
	for i, d := range GPUDevices {
		devs[i] = C.int(d)
	}
	if err := cErr(C.pir_server_create(&devs[0], C.int(len(devs)), C.uint64_t(db.DBSize), C.uint32_t(db.SlotBytes),
		C.uint32_t(GPUShardAlign), &g.srv)); err != nil {
		return nil, err
	}
	// flatten []*Slot (each a separate heap object, testing.go:12-14) in chunks of at most 64 MiB
	chunk := (64 << 20) / db.SlotBytes
	if chunk < 1 {
		chunk = 1
	}
	flat := make([]byte, chunk*db.SlotBytes)
	for lo := 0; lo < len(db.Slots); lo += chunk {
		hi := lo + chunk
		if hi > len(db.Slots) {
			hi = len(db.Slots)
		}
		for i := lo; i < hi; i++ {
			dst := flat[(i-lo)*db.SlotBytes : (i-lo+1)*db.SlotBytes]
			n := copy(dst, db.Slots[i].Data)
			for j := n; j < len(dst); j++ {
				dst[j] = 0
			}
		}
		if err := cErr(C.pir_server_upload(g.srv, C.uint64_t(lo), unsafe.Pointer(&flat[0]), C.uint64_t(hi-lo))); err != nil {
			C.pir_server_destroy(g.srv)
			return nil, err
		}
	}
	if db.Keywords != nil { // Database.SetKeywords, db.go:392-394
		kw := make([]C.uint64_t, len(db.Keywords))
		for i, k := range db.Keywords {
			kw[i] = C.uint64_t(k)
		}
		if err := cErr(C.pir_server_set_keywords(g.srv, &kw[0], C.uint64_t(len(kw)))); err != nil {
			C.pir_server_destroy(g.srv)
			return nil, err
		}
	}
	if prev, loaded := gpuTable.LoadOrStore(uintptr(unsafe.Pointer(db)), g); loaded { // another goroutine got there first
		C.pir_server_destroy(g.srv)
		return prev.(*gpuDB), nil
	}
	runtime.SetFinalizer(db, func(d *Database) { d.ReleaseGPU() })
	return g, nil
}

// ReleaseGPU frees the device copy (also run by a finalizer on the Database).
func (db *Database) ReleaseGPU() {
	if v, ok := gpuTable.LoadAndDelete(uintptr(unsafe.Pointer(db))); ok {
		g := v.(*gpuDB)
		g.mu.Lock()
		C.pir_server_destroy(g.srv)
		g.srv = nil
		g.mu.Unlock()
	}
}

// InvalidateGPU drops the device copy after db.Slots or db.Keywords were changed; the next query uploads again.
func (db *Database) InvalidateGPU() { db.ReleaseGPU() }

// marshal flattens QueryShare + dpf.Key2P + PrfKeys (query.go:14-22, dpf/common.go:13-35) into the C POD.
func marshal(q *QueryShare) (C.pir_query, error) {
	var p C.pir_query
	if !q.IsTwoParty || q.KeyTwoParty == nil {
		return p, errors.New("pir_b200: only two-party keys exist (dpf/client.go:152-156 panics for more)")
	}
	if len(q.KeyTwoParty.CW) > C.PIR_MAX_NUM_BITS || len(q.PrfKeys) > C.PIR_NUM_PRF_KEYS {
		return p, errors.New("pir_b200: key too large for the pir_query POD")
	}
	for i, k := range q.PrfKeys { // 4 keys shipped, 3 used by the 2-party PRG (dpf/common.go:62-75)
		C.memcpy(unsafe.Pointer(&p.prf_keys[i][0]), unsafe.Pointer(&k.Bytes[0]), 16)
	}
	C.memcpy(unsafe.Pointer(&p.s_init[0]), unsafe.Pointer(&q.KeyTwoParty.SInit[0]), 16)
	p.t_init = C.uint8_t(q.KeyTwoParty.TInit)
	p.share_number = C.uint8_t(q.ShareNumber)
	if q.IsKeywordBased {
		p.is_keyword = 1
	}
	p.is_two_party = 1
	p.num_bits = C.uint32_t(len(q.KeyTwoParty.CW))
	p.final_cw = C.int64_t(q.KeyTwoParty.FinalCW)
	p.group_size = C.uint32_t(q.GroupSize)
	for i, cw := range q.KeyTwoParty.CW { // 18 bytes per level (dpf/client.go:118-125)
		C.memcpy(unsafe.Pointer(&p.cw[i][0]), unsafe.Pointer(&cw[0]), 18)
	}
	return p, nil
}

func sharesFrom(out []byte, groupSize, slotBytes int) *SecretSharedQueryResult {
	shares := make([]*Slot, groupSize)
	for c := range shares { // db.go:82-89: one freshly allocated Slot per column
		shares[c] = &Slot{Data: out[c*slotBytes : (c+1)*slotBytes : (c+1)*slotBytes]}
	}
	return &SecretSharedQueryResult{SlotBytes: slotBytes, Shares: shares}
}

// PrivateSecretSharedQuery — db.go:68-72.  nprocs is accepted and ignored (the GPU count comes from GPUDevices).
func (db *Database) PrivateSecretSharedQuery(query *QueryShare, nprocs int) (*SecretSharedQueryResult, error) {
	g, err := db.gpu()
	if err != nil {
		return nil, err
	}
	p, err := marshal(query)
	if err != nil {
		return nil, err
	}
	g.mu.RLock()
	defer g.mu.RUnlock()
	if g.srv == nil {
		return nil, errReleased
	}
	if err := g.takeErr(); err != nil {
		return nil, err
	}
	out := make([]byte, query.GroupSize*db.SlotBytes)
	if err := cErr(C.pir_server_answer(g.srv, &p, unsafe.Pointer(&out[0]), C.size_t(len(out)))); err != nil {
		return nil, err
	}
	return sharesFrom(out, query.GroupSize, db.SlotBytes), nil
}

// ExpandSharedQuery — db.go:110-174: one bool per row.  The signature has no error result: a device failure is kept
// on the handle (returned by the next call that can) and nil is returned here.
func (db *Database) ExpandSharedQuery(query *QueryShare, nprocs int) []bool {
	g, err := db.gpu()
	if err != nil {
		panic(err) // no handle to keep the error on; the reference panics on a bad AES key here too (dpf/server.go:29-31)
	}
	p, err := marshal(query)
	g.mu.RLock()
	defer g.mu.RUnlock()
	if err == nil && g.srv == nil {
		err = errReleased
	}
	if err != nil {
		g.keepErr(err)
		return nil
	}
	rows := db.DBSize / query.GroupSize // db.go:114: integer division before the Ceil
	bits := make([]bool, rows)
	if rows == 0 {
		return bits
	}
	raw := make([]byte, rows)
	for s := 0; s < int(C.pir_server_num_shards(g.srv)); s++ { // every shard expands its own rows
		var sh *C.pir_db
		var first, n C.uint64_t
		if err := cErr(C.pir_server_shard(g.srv, C.int(s), &sh)); err != nil {
			g.keepErr(err)
			return nil
		}
		if err := cErr(C.pir_db_rows(sh, C.uint32_t(query.GroupSize), &first, &n)); err != nil {
			g.keepErr(err)
			return nil
		}
		if n == 0 {
			continue
		}
		if err := cErr(C.pir_expand_bits(sh, &p, (*C.uint8_t)(unsafe.Pointer(&raw[int(first)])), n)); err != nil {
			g.keepErr(err)
			return nil
		}
	}
	for i, b := range raw {
		bits[i] = b != 0
	}
	return bits
}

// PrivateSecretSharedQueryWithExpandedBits — db.go:75-107.
func (db *Database) PrivateSecretSharedQueryWithExpandedBits(query *QueryShare, bits []bool, nprocs int) (*SecretSharedQueryResult, error) {
	g, err := db.gpu()
	if err != nil {
		return nil, err
	}
	g.mu.RLock()
	defer g.mu.RUnlock()
	if g.srv == nil {
		return nil, errReleased
	}
	if err := g.takeErr(); err != nil {
		return nil, err
	}
	if bits == nil {
		return nil, errors.New("pir_b200: nil bits (an earlier ExpandSharedQuery failed)")
	}
	rowBytes := query.GroupSize * db.SlotBytes
	out := make([]byte, rowBytes)
	part := make([]byte, rowBytes)
	raw := make([]byte, len(bits))
	for i, b := range bits {
		if b {
			raw[i] = 1
		}
	}
	for s := 0; s < int(C.pir_server_num_shards(g.srv)); s++ { // XOR of the shards' partial answers (slot.go:19-30)
		var sh *C.pir_db
		var first, n C.uint64_t
		if err := cErr(C.pir_server_shard(g.srv, C.int(s), &sh)); err != nil {
			return nil, err
		}
		if err := cErr(C.pir_db_rows(sh, C.uint32_t(query.GroupSize), &first, &n)); err != nil {
			return nil, err
		}
		if n == 0 {
			continue
		}
		if int(first)+int(n) > len(raw) {
			return nil, errors.New("pir_b200: fewer bits than rows")
		}
		if err := cErr(C.pir_answer_with_bits(sh, C.uint32_t(query.GroupSize), (*C.uint8_t)(unsafe.Pointer(&raw[int(first)])), n,
			unsafe.Pointer(&part[0]), C.size_t(rowBytes))); err != nil {
			return nil, err
		}
		C.pir_xor_combine(unsafe.Pointer(&out[0]), unsafe.Pointer(&part[0]), C.size_t(rowBytes))
	}
	return sharesFrom(out, query.GroupSize, db.SlotBytes), nil
}

// PrivateSecretSharedQueryBatch answers several queries in one pass over the database (no counterpart in the
// reference, where every call is independent, db.go:68-72): BASELINE configs[3].
func (db *Database) PrivateSecretSharedQueryBatch(queries []*QueryShare) ([]*SecretSharedQueryResult, error) {
	if len(queries) == 0 {
		return nil, nil
	}
	g, err := db.gpu()
	if err != nil {
		return nil, err
	}
	ps := make([]C.pir_query, len(queries))
	for i, q := range queries {
		if ps[i], err = marshal(q); err != nil {
			return nil, err
		}
	}
	g.mu.RLock()
	defer g.mu.RUnlock()
	if g.srv == nil {
		return nil, errReleased
	}
	W := queries[0].GroupSize
	out := make([]byte, len(queries)*W*db.SlotBytes)
	if err := cErr(C.pir_server_answer_batch(g.srv, &ps[0], C.uint32_t(len(ps)), unsafe.Pointer(&out[0]), C.size_t(len(out)))); err != nil {
		return nil, err
	}
	res := make([]*SecretSharedQueryResult, len(queries))
	for i := range res {
		res[i] = sharesFrom(out[i*W*db.SlotBytes:(i+1)*W*db.SlotBytes], W, db.SlotBytes)
	}
	return res, nil
}

// MarshalBinary / UnmarshalBinary (encoding.BinaryMarshaler / BinaryUnmarshaler): the reference keeps QueryShare in
// process and has no serialisation (query.go:14-22); a client that sends its two shares to two servers uses the wire
// layout of include/pir_b200.h (100 + 18*numBits bytes, little-endian) through pir_query_encode / pir_query_decode.
func (q *QueryShare) MarshalBinary() ([]byte, error) {
	p, err := marshal(q)
	if err != nil {
		return nil, err
	}
	buf := make([]byte, int(C.pir_query_wire_size(p.num_bits)))
	if len(buf) == 0 {
		return nil, errors.New("pir_b200: numBits out of range")
	}
	var n C.size_t
	if err := cErr(C.pir_query_encode(&p, unsafe.Pointer(&buf[0]), C.size_t(len(buf)), &n)); err != nil {
		return nil, err
	}
	return buf[:int(n)], nil
}

func (q *QueryShare) UnmarshalBinary(data []byte) error {
	if len(data) == 0 {
		return errors.New("pir_b200: empty encoded share")
	}
	var p C.pir_query
	if err := cErr(C.pir_query_decode(unsafe.Pointer(&data[0]), C.size_t(len(data)), &p)); err != nil {
		return err
	}
	nb := int(p.num_bits)
	key := &dpf.Key2P{SInit: C.GoBytes(unsafe.Pointer(&p.s_init[0]), 16), TInit: byte(p.t_init), FinalCW: int(p.final_cw)}
	key.CW = make([][]byte, nb)
	for i := range key.CW {
		key.CW[i] = C.GoBytes(unsafe.Pointer(&p.cw[i][0]), 18)
	}
	q.PrfKeys = make([]*dpf.PrfKey, C.PIR_NUM_PRF_KEYS)
	for i := range q.PrfKeys {
		q.PrfKeys[i] = &dpf.PrfKey{Bytes: C.GoBytes(unsafe.Pointer(&p.prf_keys[i][0]), 16)}
	}
	q.KeyTwoParty, q.KeyMultiParty = key, nil
	q.IsKeywordBased, q.IsTwoParty = p.is_keyword != 0, p.is_two_party != 0
	q.ShareNumber, q.GroupSize = uint(p.share_number), int(p.group_size)
	return nil
}
