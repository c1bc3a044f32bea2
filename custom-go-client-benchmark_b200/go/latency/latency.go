// Package latency replaces the sort-and-index tail of benchmark-script/ssd_test/main.go (lines 146-163) and
// the histogram of README.md:21-32 with calls into libb200crc.so (C ABI: include/b200crc.h): the samples are
// copied to a B200 once and reduced there without sorting.
//
// NOT COMPILED IN THIS REPOSITORY'S CI: neither the build container nor the GPU box has a Go toolchain.  The
// same C entry points are exercised by the Python mirror (latency.py) and its GPU tests.
package latency

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../.. -lb200crc -Wl,-rpath,${SRCDIR}/../..
#include "b200crc.h"
*/
import "C"

import (
	"errors"
	"fmt"
	"sync"
	"unsafe"
)

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

// upload copies int64 samples to a device buffer the caller frees.
func upload(c *C.b200_ctx, samples []int64) (unsafe.Pointer, error) {
	var d unsafe.Pointer
	n := C.size_t(len(samples) * 8)
	if err := status(C.b200_alloc_device(c, C.int(device), n, &d)); err != nil {
		return nil, err
	}
	if err := status(C.b200_memcpy_h2d(c, C.int(device), d, unsafe.Pointer(&samples[0]), n)); err != nil {
		C.b200_free_device(c, C.int(device), d)
		return nil, err
	}
	return d, nil
}

// Percentiles holds what ssd_test prints (main.go:157-163), in the samples' unit.
type Percentiles struct {
	Avg, P20, P50, P90, P99, Min, Max int64
}

// ErrNoSamples stands for the panic the reference runs into at main.go:157 (division by len(readTime) == 0).
var ErrNoSamples = errors.New("latency: no samples")

// ComputePercentiles is main.go:146-163 without the sort: avg = sum/int64(size) (wrapping int64 sum, truncating
// division), P20 = sorted[size/5], P50 = sorted[size/2], P90 = sorted[(9*size)/10], p99 = sorted[(99*size)/100],
// min, max.  readTime is not modified (the reference sorts it in place).
func ComputePercentiles(readTime []int64) (Percentiles, error) {
	size := len(readTime)
	if size == 0 {
		return Percentiles{}, ErrNoSamples
	}
	c, err := context()
	if err != nil {
		return Percentiles{}, err
	}
	d, err := upload(c, readTime)
	if err != nil {
		return Percentiles{}, err
	}
	defer C.b200_free_device(c, C.int(device), d)
	k := [4]C.uint64_t{C.uint64_t(size / 5), C.uint64_t(size / 2), C.uint64_t((9 * size) / 10), C.uint64_t((99 * size) / 100)}
	var out [4]C.int64_t
	var st C.b200_lat_stats
	if err := status(C.b200_select_kth_avg(c, C.int(device), (*C.int64_t)(d), C.size_t(size), 1, &k[0], 4, &out[0], &st)); err != nil {
		return Percentiles{}, err
	}
	sum := int64(st.sum) // the low 64 bits: what Go's `sum += readTime[i]` holds after wrapping
	return Percentiles{Avg: sum / int64(size), P20: int64(out[0]), P50: int64(out[1]), P90: int64(out[2]),
		P99: int64(out[3]), Min: int64(st.min), Max: int64(st.max)}, nil
}

// ReadmeHistogram is numpy.histogram(x, range(20, 100, 5)) of README.md:21-23,32 over per-read durations in
// nanoseconds, each taken as Duration.Milliseconds() first (main.go:146): 15 bins [20,25) ... [90,95], and the
// number of samples outside [20, 95], which the script silently drops.  Integer milliseconds: on fractional
// Duration strings the script bins the fractions (95.3 ms is dropped there, counted in [90, 95] here); the C ABI's
// b200_parse_latency_text + b200_hist_fixed_f64 reproduce that form.
func ReadmeHistogram(durationsNs []int64) (counts [15]uint64, dropped uint64, err error) {
	if len(durationsNs) == 0 {
		return
	}
	c, err := context()
	if err != nil {
		return
	}
	d, err := upload(c, durationsNs)
	if err != nil {
		return
	}
	defer C.b200_free_device(c, C.int(device), d)
	var edges [16]C.double
	for i := range edges {
		edges[i] = C.double(20 + 5*i)
	}
	var cc [15]C.uint64_t
	var st C.b200_lat_stats
	err = status(C.b200_hist_fixed(c, C.int(device), (*C.int64_t)(d), C.size_t(len(durationsNs)), 1000000, &edges[0], 16,
		C.B200_HIST_NUMPY, &cc[0], &st))
	for i := range counts {
		counts[i] = uint64(cc[i])
	}
	return counts, uint64(st.dropped_below) + uint64(st.dropped_above), err
}
