// Package crc32c is a drop-in for the part of hash/crc32 that cloud.google.com/go/storage's Reader uses
// (crc32.New / Update / Checksum / MakeTable with the Castagnoli polynomial), computed on an NVIDIA B200 through
// libb200crc.so (C ABI: include/b200crc.h), plus the batched VerifyObjects and Combine.
//
// The hash/crc32 names keep their stdlib signatures, so a file that only uses the Castagnoli table switches over
// by changing one import line:
//
//	import crc32 "github.com/raj-prince/custom-go-client-benchmark/b200/crc32c"
//	...
//	crc32cTable := crc32.MakeTable(crc32.Castagnoli)
//	gotCRC = crc32.Update(gotCRC, crc32cTable, p[:n])
//
// As in the standard library these forms cannot return an error; a CUDA failure panics (exactly as Write does).
// UpdateErr / ChecksumErr / NewErr are the same operations with an error result for callers that want one.
//
// NOT COMPILED IN THIS REPOSITORY'S CI: neither the build container nor the GPU box has a Go toolchain.
// The file is the binding a maintainer adds; the same C entry points are exercised by the Python mirror
// and its GPU tests.
//
// cgo notes: every exported C function selects its CUDA device itself, because goroutines migrate between
// OS threads.  [][]byte cannot cross cgo directly (Go pointers to Go pointers), so VerifyObjects pins the
// slices with runtime.Pinner (Go >= 1.21, matches go.mod:3 of the reference) and passes a C array of
// pointers.  For zero-copy use AllocPinned: the library owns that memory and Go only views it.
package crc32c

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../.. -lb200crc -Wl,-rpath,${SRCDIR}/../..
#include <stdlib.h>
#include "b200crc.h"
*/
import "C"

import (
	"encoding/base64"
	"encoding/binary"
	"errors"
	"fmt"
	"hash"
	"io"
	"runtime"
	"strings"
	"sync"
	"unsafe"
)

// Castagnoli is crc32.Castagnoli.
const Castagnoli = 0x82f63b78

// Size is crc32.Size.
const Size = 4

var (
	ctxOnce sync.Once
	ctx     *C.b200_ctx
	ctxErr  error
	device  = 0
)

func status(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("b200crc: %s: %s", C.GoString(C.b200_strerror(rc)), C.GoString(C.b200_last_error()))
}

func context() (*C.b200_ctx, error) {
	ctxOnce.Do(func() {
		dev := C.int(device)
		ctxErr = status(C.b200_init(1, &dev, &ctx))
	})
	return ctx, ctxErr
}

// Table is crc32.Table for the one polynomial this package computes.  It carries no data: the tables live on
// the GPU.
type Table struct{ poly uint32 }

var castagnoliTable = &Table{poly: Castagnoli}

// MakeTable is crc32.MakeTable.  Only crc32.Castagnoli is available; any other polynomial panics, because a
// silently different checksum would be worse.
func MakeTable(poly uint32) *Table {
	if poly != Castagnoli {
		panic(fmt.Sprintf("b200crc: only the Castagnoli polynomial (%#x) is computed on the GPU, not %#x", Castagnoli, poly))
	}
	return castagnoliTable
}

func checkTable(tab *Table) {
	if tab == nil || tab.poly != Castagnoli {
		panic("b200crc: table must come from MakeTable(Castagnoli)")
	}
}

// Update is crc32.Update(crc, tab, p): Update(Update(0, tab, a), tab, b) == Update(0, tab, a+b).
func Update(crc uint32, tab *Table, p []byte) uint32 {
	checkTable(tab)
	out, err := UpdateErr(crc, p)
	if err != nil {
		panic(err)
	}
	return out
}

// Checksum is crc32.Checksum(data, tab).
func Checksum(data []byte, tab *Table) uint32 { return Update(0, tab, data) }

// New is crc32.New(tab): a hash.Hash32 computing CRC-32C.
func New(tab *Table) hash.Hash32 {
	checkTable(tab)
	h, err := NewErr()
	if err != nil {
		panic(err)
	}
	return h
}

// UpdateErr is Update with the CUDA status as an error instead of a panic.
func UpdateErr(crc uint32, p []byte) (uint32, error) {
	c, err := context()
	if err != nil {
		return 0, err
	}
	var out C.uint32_t
	var ptr unsafe.Pointer
	if len(p) > 0 {
		ptr = unsafe.Pointer(&p[0])
	}
	err = status(C.b200_crc32c_update_host(c, C.int(device), C.uint32_t(crc), ptr, C.size_t(len(p)), &out))
	return uint32(out), err
}

// ChecksumErr is Checksum with an error result.
func ChecksumErr(p []byte) (uint32, error) { return UpdateErr(0, p) }

// Combine returns the CRC of A||B given crc(A), crc(B) and len(B).  Pure host arithmetic.
func Combine(crc1, crc2 uint32, len2 int64) uint32 {
	return uint32(C.b200_crc32c_combine(C.uint32_t(crc1), C.uint32_t(crc2), C.uint64_t(len2)))
}

type digest struct{ h *C.b200_hasher }

// NewErr is New with an error result.
func NewErr() (hash.Hash32, error) {
	c, err := context()
	if err != nil {
		return nil, err
	}
	d := &digest{}
	if err := status(C.b200_hasher_new(c, C.int(device), &d.h)); err != nil {
		return nil, err
	}
	runtime.SetFinalizer(d, func(d *digest) { C.b200_hasher_free(d.h) })
	return d, nil
}

func (d *digest) Write(p []byte) (int, error) {
	if len(p) == 0 {
		return 0, nil
	}
	// hash.Hash: "It never returns an error."  A CUDA failure here is unrecoverable for the caller.
	if err := status(C.b200_hasher_write(d.h, unsafe.Pointer(&p[0]), C.size_t(len(p)))); err != nil {
		panic(err)
	}
	return len(p), nil
}

// ReadFrom implements io.ReaderFrom, so io.Copy(h, rc) replaces io.CopyBuffer(io.Discard, rc, buf)
// (main.go:140): the transport reads straight into the hasher's pinned staging buffer.
func (d *digest) ReadFrom(r io.Reader) (int64, error) {
	var total int64
	for {
		var p unsafe.Pointer
		var avail C.size_t
		if err := status(C.b200_hasher_reserve(d.h, &p, &avail)); err != nil {
			return total, err
		}
		n, err := r.Read(unsafe.Slice((*byte)(p), int(avail)))
		if n > 0 {
			if e := status(C.b200_hasher_commit(d.h, C.size_t(n))); e != nil {
				return total, e
			}
			total += int64(n)
		}
		if err == io.EOF {
			return total, nil
		}
		if err != nil {
			return total, err
		}
	}
}

func (d *digest) Sum32() uint32 {
	var out C.uint32_t
	if err := status(C.b200_hasher_sum32(d.h, &out)); err != nil {
		panic(err)
	}
	return uint32(out)
}

func (d *digest) Sum(b []byte) []byte {
	s := d.Sum32()
	return append(b, byte(s>>24), byte(s>>16), byte(s>>8), byte(s))
}

func (d *digest) Reset()         { C.b200_hasher_reset(d.h) }
func (d *digest) Size() int      { return Size }
func (d *digest) BlockSize() int { return 1 }

// ErrBadCRC mirrors storage.Reader's "storage: bad CRC on read: got %d, want %d".
type ErrBadCRC struct {
	Index     int
	Got, Want uint32
}

func (e *ErrBadCRC) Error() string {
	return fmt.Sprintf("storage: bad CRC on read: got %d, want %d (object %d)", e.Got, e.Want, e.Index)
}

// VerifyObjects checks every buffer against its expected CRC-32C in one batched GPU pass.  It returns the
// indices of the mismatching objects; a mismatch never aborts the batch.
func VerifyObjects(bufs [][]byte, want []uint32) (bad []int, err error) {
	if len(bufs) != len(want) {
		return nil, errors.New("b200crc: len(bufs) != len(want)")
	}
	c, err := context()
	if err != nil {
		return nil, err
	}
	n := len(bufs)
	if n == 0 {
		return nil, nil
	}
	ptrs := (*[1 << 30]unsafe.Pointer)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))[:n:n]
	defer C.free(unsafe.Pointer(&ptrs[0]))
	lens := make([]C.uint64_t, n)
	var pin runtime.Pinner
	defer pin.Unpin()
	for i, b := range bufs {
		lens[i] = C.uint64_t(len(b))
		if len(b) > 0 {
			pin.Pin(&b[0])
			ptrs[i] = unsafe.Pointer(&b[0])
		} else {
			ptrs[i] = nil
		}
	}
	got := make([]C.uint32_t, n)
	ok := make([]C.uint8_t, n)
	var nbad C.uint64_t
	rc := C.b200_crc32c_batch_host(c, C.int(device), (*unsafe.Pointer)(unsafe.Pointer(&ptrs[0])), &lens[0], C.size_t(n),
		(*C.uint32_t)(unsafe.Pointer(&want[0])), &got[0], &ok[0], &nbad)
	if err := status(rc); err != nil {
		return nil, err
	}
	for i := range ok {
		if ok[i] == 0 {
			bad = append(bad, i)
		}
	}
	return bad, nil
}

// AllocPinned returns a page-locked buffer owned by the library (the replacement for ReadObject's
// `buf := make([]byte, 2*MB)`, main.go:125): H2D copies from it are direct DMA.
func AllocPinned(n int) ([]byte, error) {
	c, err := context()
	if err != nil {
		return nil, err
	}
	var p unsafe.Pointer
	if err := status(C.b200_alloc_pinned(c, C.size_t(n), &p)); err != nil {
		return nil, err
	}
	return unsafe.Slice((*byte)(p), n), nil
}

// FreePinned releases a buffer obtained from AllocPinned.
func FreePinned(b []byte) error {
	c, err := context()
	if err != nil {
		return err
	}
	return status(C.b200_free_pinned(c, unsafe.Pointer(&b[0])))
}

// DecodeXGoogHash extracts the CRC-32C from an X-Goog-Hash header value ("crc32c=AAAAAA==,md5=...").
func DecodeXGoogHash(header string) (uint32, bool) {
	for _, part := range strings.Split(header, ",") {
		part = strings.TrimSpace(part)
		if v, ok := strings.CutPrefix(part, "crc32c="); ok {
			raw, err := base64.StdEncoding.DecodeString(v)
			if err != nil || len(raw) != 4 {
				return 0, false
			}
			return binary.BigEndian.Uint32(raw), true
		}
	}
	return 0, false
}
