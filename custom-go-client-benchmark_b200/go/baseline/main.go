// Stdlib-only CPU baseline for the CRC-32C verify path: times Go's own hash/crc32 (Castagnoli; SSE4.2 on
// amd64) over the BASELINE config-2 workload, on one goroutine and on GOMAXPROCS goroutines.
// It needs no modules, so it builds offline wherever a Go toolchain exists:  go run ./baseline
// NOT RUN in this repository's measurements: Go is installed neither in the build container nor on the GPU
// box; bench.py times the C restatement of the same algorithm (oracle/oracle.c) instead and says so.
package main

import (
	"fmt"
	"hash/crc32"
	"runtime"
	"sync"
	"time"
)

const (
	nobj = 4096
	size = 2 << 20
	seed = 0xC0FFEE
)

func splitmix64(seed, index uint64) uint64 {
	z := seed + (index+1)*0x9E3779B97F4A7C15
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9
	z = (z ^ (z >> 27)) * 0x94D049BB133111EB
	return z ^ (z >> 31)
}

func main() {
	tab := crc32.MakeTable(crc32.Castagnoli)
	arena := make([]byte, nobj*size)
	for i := 0; i < len(arena)/8; i++ {
		z := splitmix64(seed, uint64(i))
		for b := 0; b < 8; b++ {
			arena[8*i+b] = byte(z >> (8 * b))
		}
	}
	run := func(workers int) float64 {
		start := time.Now()
		var wg sync.WaitGroup
		for w := 0; w < workers; w++ {
			wg.Add(1)
			go func(w int) {
				defer wg.Done()
				for i := w; i < nobj; i += workers {
					_ = crc32.Update(0, tab, arena[i*size:(i+1)*size])
				}
			}(w)
		}
		wg.Wait()
		return float64(len(arena)) / time.Since(start).Seconds() / 1e9
	}
	fmt.Printf("object0 crc32c = %#08x\n", crc32.Checksum(arena[:size], tab))
	fmt.Printf("1 goroutine : %.2f GB/s\n", run(1))
	n := runtime.GOMAXPROCS(0)
	fmt.Printf("%d goroutines: %.2f GB/s\n", n, run(n))
}
