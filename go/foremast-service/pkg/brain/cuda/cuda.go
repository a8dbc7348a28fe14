// Package cuda binds libforemast_cuda.so (include/foremast_cuda.h) for foremast-service.
//
// It is pure marshalling: every number is computed by the library.  The job loop
// (package brain, one directory up in a real integration) maps verdicts to the
// document statuses of pkg/converter/converter.go:11-30 and writes reason / anomalyInfo.
//
// NOT COMPILED IN THIS REPOSITORY'S CI: the build image has no Go toolchain.  The
// Python mirror foremast_1_b200/_ffi.py binds the same symbols and is what the tests run.
//
// Build: CGO_ENABLED=1 (foremast-service/Dockerfile:12 sets 0 today), -lforemast_cuda on the
// linker path, include/foremast_cuda.h on the include path.
package cuda

/*
#cgo LDFLAGS: -lforemast_cuda
#include <stdlib.h>
#include "foremast_cuda.h"
*/
import "C"

import (
	"bytes"
	"fmt"
	"runtime"
	"unsafe"
)

// Verdicts (FM_HEALTHY / FM_UNHEALTHY / FM_UNKNOWN).
const (
	Healthy   = 0
	Unhealthy = 1
	Unknown   = 2
)

// Strategies.
const (
	StrategyCanary     = C.FM_STRAT_CANARY     // canary: current + baseline (+ historical)
	StrategyHistorical = C.FM_STRAT_HISTORICAL // rollingUpdate / continuous / hpa
)

// SeriesRec mirrors fm_series_rec (96 bytes).
type SeriesRec struct {
	Mean, Sigma, Upper, Lower float64
	U1, PU, W, PW, H, PH      float64
	NAnom                     uint32
	FirstIdx                  uint16
	Verdict, Status           uint8
	AnomMask                  uint64
}

// Ctx is one GPU + one stream; calls on a Ctx must come from one goroutine at a time.
type Ctx struct{ h *C.fm_ctx }

func init() {
	if C.fm_sizeof(4) != C.uint32_t(unsafe.Sizeof(SeriesRec{})) {
		panic("cuda: SeriesRec does not match fm_series_rec")
	}
}

// lastError reads the library's message.  With h == nil the message is thread-local in the
// library, so callers that may need it lock the goroutine to its OS thread around the failing call
// and this read (see withThread).
func lastError(h *C.fm_ctx) string { return C.GoString(C.fm_last_error(h)) }

// withThread runs f on one OS thread: the C call and the fm_last_error(NULL) that explains its
// failure must not be separated by a goroutine migration.
func withThread(f func() error) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	return f()
}

// NewCtx binds device `device`.  There is no CPU fallback: it fails without a GPU.
func NewCtx(device int) (*Ctx, error) {
	var h *C.fm_ctx
	if err := withThread(func() error {
		if rc := C.fm_ctx_create(C.int32_t(device), &h); rc != C.FM_OK {
			return fmt.Errorf("fm_ctx_create(%d): %d: %s", device, int(rc), lastError(nil))
		}
		return nil
	}); err != nil {
		return nil, err
	}
	c := &Ctx{h}
	runtime.SetFinalizer(c, func(c *Ctx) { c.Close() })
	return c, nil
}

func (c *Ctx) Close() {
	if c.h != nil {
		C.fm_ctx_destroy(c.h)
		c.h = nil
	}
}

// TypeParams is one row of the threshold table (threshold{i} / bound{i} / min_lower_bound{i}
// of deploy/foremast/3_brain/foremast-brain.yaml:26-73).
type TypeParams struct {
	K, KLow, MinLowerBound float64
	BoundMask              uint32
}

// SetParams installs the env-derived parameters; g is filled by DefaultParams and edited by the caller.
func (c *Ctx) SetParams(g *C.fm_global_params, table []TypeParams) error {
	ct := make([]C.fm_type_params, len(table))
	for i, t := range table {
		ct[i] = C.fm_type_params{k: C.double(t.K), k_low: C.double(t.KLow),
			min_lower_bound: C.double(t.MinLowerBound), bound_mask: C.uint32_t(t.BoundMask)}
	}
	if rc := C.fm_set_params(c.h, g, &ct[0], C.uint32_t(len(ct))); rc != C.FM_OK {
		return fmt.Errorf("fm_set_params: %d: %s", int(rc), lastError(c.h))
	}
	return nil
}

// DefaultParams returns the deployed defaults (foremast-brain.yaml:24-79).
func DefaultParams() (C.fm_global_params, []TypeParams) {
	var g C.fm_global_params
	ct := make([]C.fm_type_params, 8)
	n := C.uint32_t(len(ct))
	C.fm_default_params(&g, &ct[0], &n)
	out := make([]TypeParams, int(n))
	for i := range out {
		out[i] = TypeParams{float64(ct[i].k), float64(ct[i].k_low), float64(ct[i].min_lower_bound), uint32(ct[i].bound_mask)}
	}
	return g, out
}

// Pinned is a pinned host buffer from fm_alloc_pinned, viewed as []float32.
type Pinned struct {
	p unsafe.Pointer
	F []float32
}

func AllocPinned(nFloats int) (*Pinned, error) {
	var p unsafe.Pointer
	if rc := C.fm_alloc_pinned(C.size_t(nFloats*4), &p); rc != C.FM_OK {
		return nil, fmt.Errorf("fm_alloc_pinned: %d", int(rc))
	}
	return &Pinned{p, unsafe.Slice((*float32)(p), nFloats)}, nil
}

func (b *Pinned) Free() { C.fm_free_pinned(b.p); b.p = nil; b.F = nil }

// RowPitch is the row pitch in elements for rows of n points (rows start 16-byte aligned).
func RowPitch(n int) int { return (n + 3) &^ 3 }

// Batch describes S series: series-major matrices in pinned memory, optional ragged lengths.
type Batch struct {
	Strategy            uint32
	N                   int
	Cur, Base, Hist     *Pinned // Base / Hist may be nil
	NCur, NBase, NHist  int
	CurLen, BaseLen, HistLen []uint32 // optional
	MetricType          []uint8  // optional, index into the threshold table
}

// Score judges the batch (docs/guides/design.md:35,43) and fills out[0:N].
func (c *Ctx) Score(b *Batch, out []SeriesRec) error {
	if len(out) < b.N {
		return fmt.Errorf("cuda: out has %d records, need %d", len(out), b.N)
	}
	if b.N == 0 {
		return nil // nothing to judge (and &out[0] would not exist)
	}
	cb, pin := b.cBatch()
	defer pin.Unpin()
	pin.Pin(&out[0])
	if rc := C.fm_score_batch(c.h, &cb, unsafe.Pointer(&out[0]), nil, nil, nil, nil); rc != C.FM_OK {
		return fmt.Errorf("fm_score_batch: %d: %s", int(rc), lastError(c.h))
	}
	return nil
}

// StatusOf maps a roll-up of verdicts to the document status (converter.go:17-22).
func StatusOf(verdicts []uint8) string {
	unknown := len(verdicts) > 0
	for _, v := range verdicts {
		if v == Unhealthy {
			return "completed_unhealth"
		}
		if v != Unknown {
			unknown = false
		}
	}
	if unknown {
		return "completed_unknown"
	}
	return "completed_health"
}

// DecodeQueryRange turns a raw Prometheus query_range response body into samples without going
// through encoding/json: values[i] / ts[i] for i in [start[s], start[s+1]) belong to result
// series s; NaN and +-Inf samples are dropped (fm_decode_query_range, include/foremast_cuda.h).
func DecodeQueryRange(body []byte) (ts []int64, values []float32, start []uint32, err error) {
	if len(body) == 0 {
		return nil, nil, nil, fmt.Errorf("empty query_range body")
	}
	capSamples := len(body)/8 + 1
	capSeries := bytes.Count(body, []byte(`"value`)) + 2
	ts = make([]int64, capSamples)
	values = make([]float32, capSamples)
	start = make([]uint32, capSeries)
	var nSeries C.uint32_t
	runtime.LockOSThread() // fm_last_error(NULL) below is thread-local
	defer runtime.UnlockOSThread()
	n := C.fm_decode_query_range((*C.char)(unsafe.Pointer(&body[0])), C.uint64_t(len(body)),
		(*C.int64_t)(unsafe.Pointer(&ts[0])), (*C.float)(unsafe.Pointer(&values[0])), C.uint64_t(capSamples),
		(*C.uint32_t)(unsafe.Pointer(&start[0])), C.uint32_t(capSeries), &nSeries)
	if n < 0 {
		return nil, nil, nil, fmt.Errorf("fm_decode_query_range: %d: %s", int64(n), lastError(nil))
	}
	return ts[:n], values[:n], start[:int(nSeries)+1], nil
}

// ---- device-resident rows between polls (MAX_CACHE_SIZE, foremast-brain/README.md:30) ----------

// CacheConfigure gives the ctx maxBytes of HBM for baseline / history rows that stay resident
// between polls (0 = off).
func (c *Ctx) CacheConfigure(maxBytes uint64) error {
	if rc := C.fm_cache_configure(c.h, C.uint64_t(maxBytes)); rc != C.FM_OK {
		return fmt.Errorf("fm_cache_configure: %d: %s", int(rc), lastError(c.h))
	}
	return nil
}

// CacheQuery tells which row ids are resident: the job loop skips the Prometheus fetch of the
// baseline / history of those (a job is re-judged on every poll until its endTime,
// docs/guides/design.md:43, and only its current window changes).
func (c *Ctx) CacheQuery(rowIDs []uint64) ([]bool, error) {
	out := make([]bool, len(rowIDs))
	if len(rowIDs) == 0 {
		return out, nil
	}
	res := make([]uint8, len(rowIDs))
	if rc := C.fm_cache_query(c.h, (*C.uint64_t)(unsafe.Pointer(&rowIDs[0])), C.uint64_t(len(rowIDs)),
		(*C.uint8_t)(unsafe.Pointer(&res[0]))); rc != C.FM_OK {
		return nil, fmt.Errorf("fm_cache_query: %d: %s", int(rc), lastError(c.h))
	}
	for i, r := range res {
		out[i] = r != 0
	}
	return out, nil
}

// ScoreCached is Score with one 64-bit id per series (e.g. FNV-1a of job id + metric alias): the
// baseline / history rows of resident ids are not read from the host, the others are uploaded and
// become resident.  Results are identical to Score on the same rows.
func (c *Ctx) ScoreCached(b *Batch, rowIDs []uint64, out []SeriesRec) error {
	if len(out) < b.N || len(rowIDs) < b.N {
		return fmt.Errorf("cuda: need %d records and row ids", b.N)
	}
	if b.N == 0 {
		return nil
	}
	cb, pin := b.cBatch()
	defer pin.Unpin()
	pin.Pin(&out[0])
	pin.Pin(&rowIDs[0])
	if rc := C.fm_score_batch_cached(c.h, &cb, (*C.uint64_t)(unsafe.Pointer(&rowIDs[0])), unsafe.Pointer(&out[0]),
		nil, nil, nil, nil); rc != C.FM_OK {
		return fmt.Errorf("fm_score_batch_cached: %d: %s", int(rc), lastError(c.h))
	}
	return nil
}

// cBatch marshals a Batch; the caller unpins.
func (b *Batch) cBatch() (C.fm_batch, *runtime.Pinner) {
	pin := &runtime.Pinner{}
	u32 := func(s []uint32) *C.uint32_t {
		if len(s) == 0 {
			return nil
		}
		pin.Pin(&s[0])
		return (*C.uint32_t)(unsafe.Pointer(&s[0]))
	}
	cb := C.fm_batch{struct_size: C.uint32_t(unsafe.Sizeof(C.fm_batch{})), strategy: C.uint32_t(b.Strategy),
		n_series: C.uint64_t(b.N), record_kind: C.FM_REC_FULL}
	cb.cur, cb.n_cur, cb.cur_pitch, cb.cur_len = (*C.float)(b.Cur.p), C.uint32_t(b.NCur), C.uint32_t(RowPitch(b.NCur)), u32(b.CurLen)
	if b.Base != nil {
		cb.base, cb.n_base, cb.base_pitch, cb.base_len = (*C.float)(b.Base.p), C.uint32_t(b.NBase), C.uint32_t(RowPitch(b.NBase)), u32(b.BaseLen)
	}
	if b.Hist != nil {
		cb.hist, cb.n_hist, cb.hist_pitch, cb.hist_len = (*C.float)(b.Hist.p), C.uint32_t(b.NHist), C.uint32_t(RowPitch(b.NHist)), u32(b.HistLen)
	}
	if len(b.MetricType) > 0 {
		pin.Pin(&b.MetricType[0])
		cb.metric_type = (*C.uint8_t)(unsafe.Pointer(&b.MetricType[0]))
	}
	return cb, pin
}

// ---- every GPU of the box from this one process -----------------------------------------------

// Pool owns one ctx per GPU (fm_pool): the shared-nothing brain nodes of docs/guides/design.md:37-41
// as the GPUs of one box.  Use a Pool from one goroutine at a time.
type Pool struct{ h *C.fm_pool }

// NewPool binds the listed devices (nil: every visible GPU).
func NewPool(devices []int32) (*Pool, error) {
	var h *C.fm_pool
	err := withThread(func() error {
		var p *C.int32_t
		if len(devices) > 0 {
			p = (*C.int32_t)(unsafe.Pointer(&devices[0]))
		}
		if rc := C.fm_pool_create(p, C.int32_t(len(devices)), &h); rc != C.FM_OK {
			return fmt.Errorf("fm_pool_create: %d: %s", int(rc), lastError(nil))
		}
		return nil
	})
	if err != nil {
		return nil, err
	}
	return &Pool{h}, nil
}

func (p *Pool) Close() {
	if p.h != nil {
		C.fm_pool_destroy(p.h)
		p.h = nil
	}
}

func (p *Pool) Size() int { return int(C.fm_pool_size(p.h)) }

func (p *Pool) err(what string, rc C.int32_t) error {
	return fmt.Errorf("%s: %d: %s", what, int(rc), C.GoString(C.fm_pool_last_error(p.h)))
}

// SetParams installs the parameters on every ctx of the pool.
func (p *Pool) SetParams(g *C.fm_global_params, table []TypeParams) error {
	ct := make([]C.fm_type_params, len(table))
	for i, t := range table {
		ct[i] = C.fm_type_params{k: C.double(t.K), k_low: C.double(t.KLow),
			min_lower_bound: C.double(t.MinLowerBound), bound_mask: C.uint32_t(t.BoundMask)}
	}
	if rc := C.fm_pool_set_params(p.h, g, &ct[0], C.uint32_t(len(ct))); rc != C.FM_OK {
		return p.err("fm_pool_set_params", rc)
	}
	return nil
}

// CacheConfigure gives every ctx of the pool maxBytesPerDevice of HBM for resident rows.
func (p *Pool) CacheConfigure(maxBytesPerDevice uint64) error {
	if rc := C.fm_pool_cache_configure(p.h, C.uint64_t(maxBytesPerDevice)); rc != C.FM_OK {
		return p.err("fm_pool_cache_configure", rc)
	}
	return nil
}

// Score cuts the batch into one contiguous range of series per GPU, scores the ranges
// concurrently and fills out[0:N] (rowIDs == nil: no row cache).
func (p *Pool) Score(b *Batch, rowIDs []uint64, out []SeriesRec) error {
	if len(out) < b.N || (rowIDs != nil && len(rowIDs) < b.N) {
		return fmt.Errorf("cuda: need %d records (and row ids)", b.N)
	}
	if b.N == 0 {
		return nil
	}
	cb, pin := b.cBatch()
	defer pin.Unpin()
	pin.Pin(&out[0])
	var ids *C.uint64_t
	if rowIDs != nil {
		pin.Pin(&rowIDs[0])
		ids = (*C.uint64_t)(unsafe.Pointer(&rowIDs[0]))
	}
	if rc := C.fm_pool_score_batch(p.h, &cb, ids, unsafe.Pointer(&out[0]), nil, nil, nil, nil); rc != C.FM_OK {
		return p.err("fm_pool_score_batch", rc)
	}
	return nil
}

// ---- two metrics judged jointly (bivariate normal, docs/guides/design.md:67-76) ---------------

// PairRec mirrors fm_pair_rec (64 bytes).
type PairRec struct {
	MeanX, MeanY, VarX, VarY, CovXY, D2Max float64
	NAnom                                  uint32
	FirstIdx                               uint16
	Verdict, Status                        uint8
	AnomMask                               uint64
}

// ScorePairs judges n metric pairs: rows of pair p are curX[p], curY[p] (nCur points) and histX[p],
// histY[p] (nHist points), all in pinned matrices of pitch RowPitch(n).
func (c *Ctx) ScorePairs(n int, curX, curY *Pinned, nCur int, histX, histY *Pinned, nHist int, metricType []uint8, out []PairRec) error {
	if len(out) < n {
		return fmt.Errorf("cuda: out has %d records, need %d", len(out), n)
	}
	if n == 0 {
		return nil
	}
	if C.fm_sizeof(12) != C.uint32_t(unsafe.Sizeof(PairRec{})) {
		return fmt.Errorf("cuda: PairRec does not match fm_pair_rec")
	}
	var pin runtime.Pinner
	defer pin.Unpin()
	pb := C.fm_pair_batch{struct_size: C.uint32_t(unsafe.Sizeof(C.fm_pair_batch{})), n_pairs: C.uint64_t(n)}
	pb.cur_x, pb.cur_y, pb.n_cur, pb.cur_pitch = (*C.float)(curX.p), (*C.float)(curY.p), C.uint32_t(nCur), C.uint32_t(RowPitch(nCur))
	pb.hist_x, pb.hist_y, pb.n_hist, pb.hist_pitch = (*C.float)(histX.p), (*C.float)(histY.p), C.uint32_t(nHist), C.uint32_t(RowPitch(nHist))
	if len(metricType) > 0 {
		pin.Pin(&metricType[0])
		pb.metric_type = (*C.uint8_t)(unsafe.Pointer(&metricType[0]))
	}
	pin.Pin(&out[0])
	if rc := C.fm_score_pairs(c.h, &pb, (*C.fm_pair_rec)(unsafe.Pointer(&out[0])), nil); rc != C.FM_OK {
		return fmt.Errorf("fm_score_pairs: %d: %s", int(rc), lastError(c.h))
	}
	return nil
}
