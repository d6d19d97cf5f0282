// go_vectors — pins the CPU oracle against the REAL reference, one command away.
//
// Reads tests/golden/dpf_pir_vectors.json (the cases: 97-byte random tapes, domains, points, database shapes),
// re-computes every case with the UNMODIFIED `dpf` package of the reference (github.com/sachaservan/pir/dpf: stdlib
// only) and prints a JSON document with the same keys, to be saved as tests/golden/go_vectors.json.
// tests/test_oracle_golden.py::test_go_vectors_if_present then diffs the two byte for byte.
//
// How the reference is made deterministic without touching it: crypto/rand.Reader is replaced by a reader that
// replays the case's tape.  ClientInitialize draws 4 x 16 bytes (dpf/client.go:27), GenerateTwoServer draws 17 and
// then 16 (dpf/client.go:60,67) — 97 bytes, in that order.
//
// What is NOT taken from the reference: package `pir` itself cannot be built offline (db.go:8-9 import un-vendored
// Paillier / gmp modules), so the three lines of db.go that sit between dpf and the answer are restated here and
// marked: the bit rule (db.go:142), the row XOR (db.go:91-104, slot.go:19-30) and numBits (db.go:114-121).
// The synthetic database generator is this repo's (oracle/pir_oracle.c orc_synth_word), not the reference's.
//
// No Go toolchain exists in the container this repo was built in: this file has never been compiled there.
// See README.md next to it for the two commands.
package main

import (
	"crypto/rand"
	"encoding/binary"
	"encoding/hex"
	"encoding/json"
	"fmt"
	"math"
	"os"

	"github.com/sachaservan/pir/dpf"
)

// tapeReader replays a fixed byte string as the process's randomness.
type tapeReader struct {
	buf []byte
	pos int
}

func (t *tapeReader) Read(p []byte) (int, error) {
	if t.pos+len(p) > len(t.buf) {
		panic(fmt.Sprintf("tape exhausted: want %d more bytes at offset %d of %d", len(p), t.pos, len(t.buf)))
	}
	n := copy(p, t.buf[t.pos:t.pos+len(p)])
	t.pos += n
	return n, nil
}

// keysFromTape = dpf.ClientInitialize + GenerateTwoServer with crypto/rand replaying `tape`.
func keysFromTape(tape []byte, numBits uint, a, b uint) (*dpf.Dpf, []*dpf.Key2P) {
	tr := &tapeReader{buf: tape}
	rand.Reader = tr
	f := dpf.ClientInitialize(numBits)
	keys := f.GenerateTwoServer(a, b)
	if tr.pos != 97 {
		panic(fmt.Sprintf("the reference drew %d random bytes, expected 97", tr.pos))
	}
	return f, keys
}

// marshalQuery lays one server's share out as the `pir_query` POD of include/pir_b200.h (1256 bytes, little endian).
func marshalQuery(f *dpf.Dpf, k *dpf.Key2P, share int, numBits uint, groupSize uint32, isKeyword bool) []byte {
	out := make([]byte, 1256)
	for i := 0; i < 4; i++ {
		copy(out[16*i:16*i+16], f.PrfKeys[i].Bytes)
	}
	copy(out[64:80], k.SInit)
	out[80] = k.TInit
	out[81] = byte(share)
	if isKeyword {
		out[82] = 1
	}
	out[83] = 1 // IsTwoParty
	binary.LittleEndian.PutUint32(out[84:88], uint32(numBits))
	binary.LittleEndian.PutUint64(out[88:96], uint64(int64(k.FinalCW)))
	binary.LittleEndian.PutUint32(out[96:100], groupSize)
	// out[100:104] reserved = 0
	for i := range k.CW {
		copy(out[104+18*i:104+18*i+18], k.CW[i])
	}
	return out
}

// ---- this repo's synthetic database (oracle/pir_oracle.c orc_synth_word / orc_synth_fill) ----
func synthWord(seed, slot uint64, w uint32) uint64 {
	z := seed + slot*0x9E3779B97F4A7C15 + uint64(w)*0xD1B54A32D192ED03
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9
	z = (z ^ (z >> 27)) * 0x94D049BB133111EB
	return z ^ (z >> 31)
}

func synthSlot(seed, slot uint64, slotBytes int) []byte {
	out := make([]byte, slotBytes)
	for j := 0; j < slotBytes; j++ {
		out[j] = byte(synthWord(seed, slot, uint32(j/8)) >> (8 * uint(j%8)))
	}
	return out
}

func packBitsLE(bits []bool) []byte {
	out := make([]byte, (len(bits)+7)/8)
	for i, b := range bits {
		if b {
			out[i/8] |= 1 << uint(i%8)
		}
	}
	return out
}

func num(m map[string]interface{}, key string) uint64 {
	return uint64(m[key].(float64))
}

func pirCase(c map[string]interface{}) map[string]interface{} {
	tape, _ := hex.DecodeString(c["tape"].(string))
	dbSize, slotBytes, W := int(num(c, "db_size")), int(num(c, "slot_bytes")), int(num(c, "group_size"))
	index, seed := uint(num(c, "index")), num(c, "db_seed")
	// db.go:114-121 (restated): dimHeight = DBSize / GroupSize (integer division), numBits = floor(log2 H) + 1
	dimHeight := int(math.Ceil(float64(dbSize / W)))
	numBits := uint(math.Log2(float64(dimHeight)) + 1)
	f, keys := keysFromTape(tape, numBits, index, 1) // query.go:68-74: b is always 1
	out := map[string]interface{}{}
	for k, v := range c {
		out[k] = v
	}
	out["num_bits"] = numBits
	for share := 0; share < 2; share++ {
		out[fmt.Sprintf("query%d", share)] = hex.EncodeToString(marshalQuery(f, keys[share], share, numBits, uint32(W), false))
		pf := dpf.ServerInitialize(f.PrfKeys, numBits) // db.go:124
		bits := make([]bool, dimHeight)
		for i := 0; i < dimHeight; i++ {
			res := pf.Evaluate2P(uint(share), keys[share], uint(i))  // db.go:140
			bits[i] = (int(math.Abs(float64(res%2))) == 0)           // db.go:142 (restated)
		}
		out[fmt.Sprintf("bits%d_packed_le", share)] = hex.EncodeToString(packBitsLE(bits))
		// db.go:91-104 + slot.go:19-30 (restated) over the synthetic slots
		answer := make([]byte, W*slotBytes)
		for row := 0; row < dimHeight; row++ {
			if !bits[row] {
				continue
			}
			for col := 0; col < W; col++ {
				slotIndex := row*W + col
				if slotIndex >= dbSize {
					break
				}
				s := synthSlot(seed, uint64(slotIndex), slotBytes)
				for j := 0; j < slotBytes; j++ {
					answer[col*slotBytes+j] ^= s[j]
				}
			}
		}
		out[fmt.Sprintf("answer%d", share)] = hex.EncodeToString(answer)
	}
	return out
}

func dpfCase(c map[string]interface{}) map[string]interface{} {
	tape, _ := hex.DecodeString(c["tape"].(string))
	numBits := uint(num(c, "num_bits"))
	// a and points may exceed 2^53: they are read as json.Number below
	a := uint(c["a_u64"].(uint64))
	b := uint(num(c, "b"))
	f, keys := keysFromTape(tape, numBits, a, b)
	out := map[string]interface{}{}
	for k, v := range c {
		if k != "a_u64" && k != "points_u64" {
			out[k] = v
		}
	}
	points := c["points_u64"].([]uint64)
	for share := 0; share < 2; share++ {
		out[fmt.Sprintf("query%d", share)] = hex.EncodeToString(marshalQuery(f, keys[share], share, numBits, 1, false))
		pf := dpf.ServerInitialize(f.PrfKeys, numBits)
		vals := make([]int64, len(points))
		for i, x := range points {
			vals[i] = int64(pf.Evaluate2P(uint(share), keys[share], uint(x)))
		}
		out[fmt.Sprintf("values%d", share)] = vals
	}
	return out
}

func main() {
	if len(os.Args) != 2 {
		fmt.Fprintln(os.Stderr, "usage: go_vectors <path to tests/golden/dpf_pir_vectors.json>  > go_vectors.json")
		os.Exit(2)
	}
	fh, err := os.Open(os.Args[1])
	if err != nil {
		panic(err)
	}
	defer fh.Close()
	dec := json.NewDecoder(fh)
	dec.UseNumber() // 64-bit points do not fit a float64
	var doc map[string]interface{}
	if err := dec.Decode(&doc); err != nil {
		panic(err)
	}
	toFloat := func(m map[string]interface{}) map[string]interface{} {
		out := map[string]interface{}{}
		for k, v := range m {
			if n, ok := v.(json.Number); ok {
				if u, err := n.Int64(); err == nil && u >= 0 && u < (1<<53) {
					out[k] = float64(u)
				} else {
					out[k] = n
				}
			} else {
				out[k] = v
			}
		}
		return out
	}
	res := map[string]interface{}{
		"about": "computed by tools/go_vectors with the unmodified reference dpf package; same schema as dpf_pir_vectors.json",
	}
	var pirOut, dpfOut []interface{}
	for _, raw := range doc["pir"].([]interface{}) {
		pirOut = append(pirOut, pirCase(toFloat(raw.(map[string]interface{}))))
	}
	for _, raw := range doc["dpf"].([]interface{}) {
		m := raw.(map[string]interface{})
		c := toFloat(m)
		au, err := m["a"].(json.Number).Int64()
		if err != nil { // above 2^63 - 1
			var u uint64
			fmt.Sscan(m["a"].(json.Number).String(), &u)
			c["a_u64"] = u
		} else {
			c["a_u64"] = uint64(au)
		}
		c["a"] = m["a"]
		var pts []uint64
		for _, p := range m["points"].([]interface{}) {
			var u uint64
			fmt.Sscan(p.(json.Number).String(), &u)
			pts = append(pts, u)
		}
		c["points_u64"] = pts
		c["points"] = m["points"]
		dpfOut = append(dpfOut, dpfCase(c))
	}
	res["pir"] = pirOut
	res["dpf"] = dpfOut
	enc := json.NewEncoder(os.Stdout)
	enc.SetIndent("", " ")
	if err := enc.Encode(res); err != nil {
		panic(err)
	}
}
