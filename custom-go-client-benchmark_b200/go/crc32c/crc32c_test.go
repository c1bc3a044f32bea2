// Parity tests of the cgo package against Go's own hash/crc32 (the implementation storage.Reader uses).
// NOT RUN in this repository: no Go toolchain in the build image or on the GPU box.  They restate, for Go,
// what tests/test_crc_gpu.py checks through the Python mirror against the C oracle.
//
//	cd custom-go-client-benchmark_b200/go && go test ./crc32c      (needs a B200 and libb200crc.so)
package crc32c

import (
	"bytes"
	"hash/crc32"
	"io"
	"math/rand"
	"testing"
)

var castagnoli = crc32.MakeTable(crc32.Castagnoli)

func payload(n int, seed int64) []byte {
	b := make([]byte, n)
	rand.New(rand.NewSource(seed)).Read(b)
	return b
}

func TestKnownAnswers(t *testing.T) {
	for _, c := range []struct {
		in   string
		want uint32
	}{{"", 0}, {"a", 0xC1D04330}, {"123456789", 0xE3069283}} {
		got, err := ChecksumErr([]byte(c.in))
		if err != nil || got != c.want {
			t.Fatalf("Checksum(%q) = %#x, %v; want %#x", c.in, got, err, c.want)
		}
	}
}

func TestUpdateMatchesStdlibOnAnySplit(t *testing.T) {
	data := payload(5<<20+123, 1)
	want := crc32.Update(0, castagnoli, data)
	rng := rand.New(rand.NewSource(2))
	var crc uint32
	for off := 0; off < len(data); {
		n := 1 + rng.Intn(700000)
		if off+n > len(data) {
			n = len(data) - off
		}
		var err error
		if crc, err = UpdateErr(crc, data[off:off+n]); err != nil {
			t.Fatal(err)
		}
		off += n
	}
	if crc != want {
		t.Fatalf("chunked Update = %#x, want %#x", crc, want)
	}
}

func TestHash32Interface(t *testing.T) {
	h, err := NewErr()
	if err != nil {
		t.Fatal(err)
	}
	ref := crc32.New(castagnoli)
	if h.Size() != ref.Size() || h.BlockSize() != ref.BlockSize() {
		t.Fatal("Size/BlockSize differ from hash/crc32")
	}
	data := payload(9<<20, 3) // crosses the 8 MiB staging buffer
	for off := 0; off < len(data); off += 2 << 20 { // the read loop's buffer size, main.go:125
		h.Write(data[off : off+2<<20])
		ref.Write(data[off : off+2<<20])
		if h.Sum32() != ref.Sum32() {
			t.Fatalf("Sum32 differs after %d bytes", off+2<<20)
		}
	}
	if !bytes.Equal(h.Sum([]byte{1, 2}), ref.Sum([]byte{1, 2})) {
		t.Fatal("Sum (big-endian append) differs")
	}
	h.Reset()
	ref.Reset()
	if n, err := io.Copy(h, bytes.NewReader(data)); err != nil || n != int64(len(data)) { // uses ReadFrom
		t.Fatal(n, err)
	}
	ref.Write(data)
	if h.Sum32() != ref.Sum32() {
		t.Fatal("io.Copy through ReadFrom differs")
	}
}

func TestVerifyObjectsFlagsExactlyTheCorruptObject(t *testing.T) {
	const n, size = 48, 2 << 20 // 48 workers x 2 MiB (main.go:36,125)
	bufs := make([][]byte, n)
	want := make([]uint32, n)
	for i := range bufs {
		bufs[i] = payload(size, int64(100+i))
		want[i] = crc32.Checksum(bufs[i], castagnoli)
	}
	if bad, err := VerifyObjects(bufs, want); err != nil || len(bad) != 0 {
		t.Fatal(bad, err)
	}
	bufs[17][777] ^= 1
	bad, err := VerifyObjects(bufs, want)
	if err != nil || len(bad) != 1 || bad[0] != 17 {
		t.Fatalf("bad = %v, err = %v; want [17]", bad, err)
	}
}

func TestCombine(t *testing.T) {
	a, b := payload(12345, 7), payload(67890, 8)
	ca, cb := crc32.Checksum(a, castagnoli), crc32.Checksum(b, castagnoli)
	if got := Combine(ca, cb, int64(len(b))); got != crc32.Checksum(append(a, b...), castagnoli) {
		t.Fatalf("Combine = %#x", got)
	}
}

func TestXGoogHash(t *testing.T) {
	if v, ok := DecodeXGoogHash("md5=XrY7u+Ae7tCTyyK7j1rNww==, crc32c=4waSgw=="); !ok || v != 0xE3069283 {
		t.Fatalf("DecodeXGoogHash = %#x, %v", v, ok)
	}
}

// The stdlib-shaped forms: the same expressions a caller of hash/crc32 writes, with this package under the
// import alias crc32 (here: unqualified), must give what hash/crc32 gives.
func TestStdlibSignatures(t *testing.T) {
	tab := MakeTable(Castagnoli)
	data := payload(3<<20+17, 7)
	if got, want := Checksum(data, tab), crc32.Checksum(data, castagnoli); got != want {
		t.Fatalf("Checksum = %#x, want %#x", got, want)
	}
	if got, want := Update(Update(0, tab, data[:1000]), tab, data[1000:]), crc32.Checksum(data, castagnoli); got != want {
		t.Fatalf("Update chain = %#x, want %#x", got, want)
	}
	var h hash32 = New(tab)
	h.Write(data)
	if got, want := h.Sum32(), crc32.Checksum(data, castagnoli); got != want {
		t.Fatalf("New(tab).Sum32 = %#x, want %#x", got, want)
	}
	defer func() {
		if recover() == nil {
			t.Fatal("MakeTable(IEEE) must panic: only Castagnoli is computed")
		}
	}()
	MakeTable(crc32.IEEE)
}

type hash32 interface {
	io.Writer
	Sum32() uint32
}
