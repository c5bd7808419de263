package main

// GPU-backed replacement for backend/go/stores/store.go.
//
// Same type name, same embedded mixin idea, same methods and error texts as the
// reference Store; every loop over keys happens behind cgo in libtfpstore.so
// (include/tfp_store.h).  Values stay in Go: the library hands out int64 ids and
// this file keeps id -> []byte.  The reference's main.go is kept as it is
// (backend/go/stores/main.go:18-26: flag --addr, grpc.StartServer(addr, NewStore())).
//
// What this host adds over the reference, all through the unchanged gRPC surface:
//   - the store spans every GPU named in TFP_DEVICES ("0,1,2,3" or "all"): the loader hands a
//     backend only --addr and its environment (pkg/model/process.go:112-113);
//   - concurrent StoresFind calls are drained by one goroutine and served by ONE
//     tfp_store_find_batch (3 or more queries ride the tensor cores);
//   - Load(opts) restores a snapshot from opts.ModelFile (the loader already passes it,
//     pkg/model/initializers.go:397-404) and SIGTERM saves one there;
//   - Status lists HBM bytes per GPU in the memory breakdown (pkg/grpc/base/base.go:77-81).
//
// NOT COMPILED in the build image (no Go toolchain, no generated pb.go): the Python host
// localai_tfp_b200/store.py + server.py is the tested mirror of this file, call for call.

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../localai_tfp_b200 -ltfpstore
#include <stdlib.h>
#include "tfp_store.h"
*/
import "C"

import (
	"bufio"
	"crypto/rand"
	"encoding/binary"
	"errors"
	"fmt"
	"io"
	"os"
	"os/signal"
	"path/filepath"
	"runtime"
	"strconv"
	"strings"
	"sync"
	"syscall"
	"unsafe"

	"github.com/mudler/LocalAI/pkg/grpc/base"
	pb "github.com/mudler/LocalAI/pkg/grpc/proto"
)

type findReq struct {
	key  []float32
	topk int
	done chan findRes
}
type findRes struct {
	res pb.StoresFindResult
	err error
}

type Store struct {
	// base.Base (Locking() == false) instead of the reference's base.SingleThread:
	// the library serialises mutations itself and lets finds overlap.
	base.Base

	mu       sync.RWMutex // guards the values map, for the instant a lookup or an update takes
	setMu    sync.Mutex   // one mutation at a time, from its library call until its values are in the map
	h        *C.tfp_store
	values   map[int64][]byte
	finds    chan findReq
	snapshot string // opts.ModelFile of Load, "" = no persistence
}

// devices reads TFP_DEVICES: "" -> GPU 0, "all" -> every visible GPU, "0,2,3" -> those.
func devices() []C.int {
	spec := strings.TrimSpace(os.Getenv("TFP_DEVICES"))
	if spec == "" {
		return []C.int{0}
	}
	if strings.EqualFold(spec, "all") {
		n := 0
		for ; n < 8; n++ { // /proc/driver/nvidia/gpus has one entry per GPU; CUDA_VISIBLE_DEVICES narrows it
			if _, err := os.Stat(fmt.Sprintf("/dev/nvidia%d", n)); err != nil {
				break
			}
		}
		if v := os.Getenv("CUDA_VISIBLE_DEVICES"); v != "" {
			n = len(strings.Split(v, ","))
		}
		if n < 1 {
			n = 1
		}
		out := make([]C.int, n)
		for i := range out {
			out[i] = C.int(i)
		}
		return out
	}
	var out []C.int
	for _, f := range strings.Split(spec, ",") {
		if v, err := strconv.Atoi(strings.TrimSpace(f)); err == nil {
			out = append(out, C.int(v))
		}
	}
	if len(out) == 0 {
		out = []C.int{0}
	}
	return out
}

func NewStore() *Store {
	s := &Store{values: map[int64][]byte{}, finds: make(chan findReq, 1024)}
	devs := devices()
	cfg := C.tfp_cfg{device: devs[0]}
	// tfp_last_error is thread-local: keep the goroutine on its OS thread between a failing call and the read
	runtime.LockOSThread()
	rc := C.tfp_store_open_devices(&cfg, &devs[0], C.int(len(devs)), &s.h)
	msg := C.GoString(C.tfp_last_error())
	runtime.UnlockOSThread()
	if rc != 0 {
		panic("tfp_store_open_devices: " + msg)
	}
	go s.batcher()
	// the watchdog stops backends with SIGTERM (pkg/model/process.go): save before leaving
	sig := make(chan os.Signal, 1)
	signal.Notify(sig, syscall.SIGTERM)
	go func() {
		<-sig
		s.setMu.Lock()
		s.mu.Lock()
		if s.snapshot != "" {
			_ = s.save(s.snapshot)
		}
		C.tfp_store_close(s.h)
		os.Exit(0)
	}()
	return s
}

// call runs one C entry point and, on failure, returns its message.  The error text is thread-local
// in the library and goroutines move between OS threads from one cgo call to the next, so the pair
// (call, tfp_last_error) runs with the goroutine pinned to its thread.
func call(f func() C.int) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if rc := f(); rc != 0 {
		return errors.New(C.GoString(C.tfp_last_error()))
	}
	return nil
}

// pack copies the request's keys into one page-locked C block (cgo forbids lending Go pointers to Go
// pointers; from pinned memory the library's copy to the GPU runs at the PCIe rate).
func pack(keys []*pb.StoresKey) (*C.float, int, func(), error) {
	if len(keys) == 0 {
		return nil, 0, func() {}, nil
	}
	dim := len(keys[0].Floats)
	for _, k := range keys {
		if len(k.Floats) != dim {
			return nil, 0, func() {}, fmt.Errorf("Try to add key with length %d when existing length is %d", len(k.Floats), dim)
		}
	}
	n := len(keys)
	p := C.tfp_host_alloc(C.size_t(n*dim*4 + 4))
	if p == nil {
		return nil, 0, func() {}, errors.New(C.GoString(C.tfp_last_error()))
	}
	buf := (*C.float)(p)
	dst := unsafe.Slice((*float32)(unsafe.Pointer(buf)), n*dim)
	for i, k := range keys {
		copy(dst[i*dim:(i+1)*dim], k.Floats)
	}
	return buf, dim, func() { C.tfp_host_free(p) }, nil
}

// Load: the reference's is a no-op (store.go:101-103).  Here opts.ModelFile, when it names an existing
// snapshot, is restored; the path is remembered for the save on SIGTERM.  A snapshot that cannot be restored
// is logged and ignored: the backend must come up (the loader retries a backend that fails LoadModel).
func (s *Store) Load(opts *pb.ModelOptions) error {
	if opts == nil || opts.ModelFile == "" {
		return nil
	}
	s.setMu.Lock()
	defer s.setMu.Unlock()
	s.mu.Lock()
	defer s.mu.Unlock()
	s.snapshot = opts.ModelFile
	if _, err := os.Stat(s.snapshot); err != nil {
		return nil
	}
	if err := s.restore(s.snapshot); err != nil {
		fmt.Fprintf(os.Stderr, "local-store: snapshot %s not restored (%v); starting empty\n", s.snapshot, err)
		s.values = map[int64][]byte{}
	}
	return nil
}

func (s *Store) StoresSet(opts *pb.StoresSetOptions) error {
	buf, dim, free, err := pack(opts.Keys)
	defer free()
	if err != nil {
		return err
	}
	n := len(opts.Keys)
	ids := make([]int64, n+1)
	// Finds are NOT held up by a StoresSet: the library appends beside them and this host takes the map's lock only
	// to insert the values.  (A find that already sees a new row but not yet its value waits for setMu: valuesOf.)
	s.setMu.Lock()
	defer s.setMu.Unlock()
	if err := call(func() C.int {
		return C.tfp_store_set(s.h, buf, C.int64_t(n), C.int(dim), C.int64_t(len(opts.Values)),
			(*C.int64_t)(unsafe.Pointer(&ids[0])), nil)
	}); err != nil {
		return err
	}
	s.mu.Lock()
	for i := 0; i < n; i++ {
		s.values[ids[i]] = opts.Values[i].Bytes
	}
	s.mu.Unlock()
	return nil
}

// valuesOf looks up the values of the ids a find returned.  An id whose value is not in the map yet belongs to a
// StoresSet that has published its rows but not finished: wait for it (setMu), then look again.
func (s *Store) valuesOf(ids []int64) ([][]byte, bool) {
	out := make([][]byte, len(ids))
	missing := false
	s.mu.RLock()
	for i, id := range ids {
		v, ok := s.values[id]
		out[i] = v
		missing = missing || !ok
	}
	s.mu.RUnlock()
	if missing {
		s.setMu.Lock()
		s.setMu.Unlock()
		missing = false
		s.mu.RLock()
		for i, id := range ids {
			v, ok := s.values[id]
			out[i] = v
			missing = missing || !ok // still missing: the row was deleted after the find saw it
		}
		s.mu.RUnlock()
	}
	return out, !missing
}

func (s *Store) StoresDelete(opts *pb.StoresDeleteOptions) error {
	buf, dim, free, err := pack(opts.Keys)
	defer free()
	if err != nil {
		return err
	}
	n := len(opts.Keys)
	ids := make([]int64, n+1)
	s.setMu.Lock()
	defer s.setMu.Unlock()
	if err := call(func() C.int {
		return C.tfp_store_delete(s.h, buf, C.int64_t(n), C.int(dim), (*C.int64_t)(unsafe.Pointer(&ids[0])))
	}); err != nil {
		return err
	}
	s.mu.Lock()
	for i := 0; i < n; i++ {
		if ids[i] >= 0 {
			delete(s.values, ids[i])
		}
	}
	s.mu.Unlock()
	return nil
}

func (s *Store) StoresGet(opts *pb.StoresGetOptions) (pb.StoresGetResult, error) {
	buf, dim, free, err := pack(opts.Keys)
	defer free()
	if err != nil {
		return pb.StoresGetResult{}, err
	}
	n := len(opts.Keys)
	ids := make([]int64, n+1)
	order := make([]int64, n+1)
	var found C.int64_t
	if err := call(func() C.int {
		return C.tfp_store_get(s.h, buf, C.int64_t(n), C.int(dim), (*C.int64_t)(unsafe.Pointer(&ids[0])),
			(*C.int64_t)(unsafe.Pointer(&order[0])), &found)
	}); err != nil {
		return pb.StoresGetResult{}, err
	}
	res := pb.StoresGetResult{}
	hit := make([]int64, 0, int(found))
	for j := 0; j < int(found); j++ {
		hit = append(hit, ids[order[j]])
	}
	vals, _ := s.valuesOf(hit)
	for j := 0; j < int(found); j++ {
		res.Keys = append(res.Keys, &pb.StoresKey{Floats: opts.Keys[order[j]].Floats})
		res.Values = append(res.Values, &pb.StoresValue{Bytes: vals[j]})
	}
	return res, nil
}

// StoresFind hands the request to the batcher and waits for its answer.
func (s *Store) StoresFind(opts *pb.StoresFindOptions) (pb.StoresFindResult, error) {
	var key []float32
	if opts.Key != nil {
		key = opts.Key.Floats
	}
	req := findReq{key: key, topk: int(opts.TopK), done: make(chan findRes, 1)}
	s.finds <- req
	r := <-req.done
	return r.res, r.err
}

// batcher: whatever has queued up while the previous batch ran is served together, grouped by (top-k, key
// length).  A lone request is served at once: nothing waits for company.  (server.py _FindBatcher.)
func (s *Store) batcher() {
	for first := range s.finds {
		group := []findReq{first}
	drain:
		for len(group) < 256 {
			select {
			case r := <-s.finds:
				group = append(group, r)
			default:
				break drain
			}
		}
		type gk struct{ topk, dim int }
		by := map[gk][]findReq{}
		for _, r := range group {
			k := gk{r.topk, len(r.key)}
			by[k] = append(by[k], r)
		}
		for k, reqs := range by {
			s.findBatch(k.topk, k.dim, reqs)
		}
	}
}

func (s *Store) findBatch(topk, dim int, reqs []findReq) { s.findBatchN(topk, dim, reqs, 0) }

func (s *Store) findBatchN(topk, dim int, reqs []findReq, attempt int) {
	nq := len(reqs)
	var st C.tfp_stats
	C.tfp_store_stats(s.h, &st)
	slots := topk
	if int64(slots) > int64(st.rows) {
		slots = int(st.rows) // a result holds min(TopK, rows) entries (store.go:386-388)
	}
	if slots < 1 {
		slots = 1
	}
	callK := topk
	if topk >= 1 {
		callK = slots
	}
	q := make([]float32, nq*dim+1)
	for i, r := range reqs {
		copy(q[i*dim:(i+1)*dim], r.key)
	}
	ns := make([]C.int, nq)
	ids := make([]int64, nq*slots)
	sims := make([]float32, nq*slots)
	keys := make([]float32, nq*slots*dim+1)
	err := call(func() C.int {
		return C.tfp_store_find_batch(s.h, (*C.float)(unsafe.Pointer(&q[0])), C.int(nq), C.int(dim), C.int(callK), &ns[0],
			(*C.int64_t)(unsafe.Pointer(&ids[0])), (*C.float)(unsafe.Pointer(&sims[0])), (*C.float)(unsafe.Pointer(&keys[0])))
	})
	for i, r := range reqs {
		if err != nil {
			r.done <- findRes{err: err} // same text for every request of the group: they share top-k and key length
			continue
		}
		n := int(ns[i])
		res := pb.StoresFindResult{Similarities: sims[i*slots : i*slots+n]}
		vals, complete := s.valuesOf(ids[i*slots : i*slots+n])
		if !complete && attempt < 3 {
			// a row deleted between the find and the look-up has no value any more: answer this request from the
			// state after the delete instead of handing out a row without its value
			s.findBatchN(topk, dim, []findReq{r}, attempt+1)
			continue
		}
		for j := 0; j < n; j++ {
			at := i*slots + j
			res.Keys = append(res.Keys, &pb.StoresKey{Floats: keys[at*dim : (at+1)*dim]})
			res.Values = append(res.Values, &pb.StoresValue{Bytes: vals[j]})
		}
		r.done <- findRes{res: res}
	}
}

// Status: the base mixin's process figures plus HBM bytes per GPU.
func (s *Store) Status() (pb.StatusResponse, error) {
	resp, err := s.Base.Status()
	if err != nil || resp.Memory == nil {
		return resp, err
	}
	n := int(C.tfp_store_shard_count(s.h))
	var rows uint64
	for g := 0; g < n; g++ {
		var st C.tfp_stats
		if C.tfp_store_shard_stats(s.h, C.int(g), &st) == 0 {
			resp.Memory.Breakdown[fmt.Sprintf("hbm-bytes-gpu%d", int(st.device))] += uint64(st.hbm_bytes)
			rows += uint64(st.rows)
		}
	}
	resp.Memory.Breakdown["rows"] = rows
	resp.State = pb.StatusResponse_READY
	return resp, nil
}

// ---- snapshot: keys / ids / flags by the library, id -> value by this host --------------------------------
// Values file `<path>.values.<tag as 16 hex digits>`, little endian:
//     "TFPVALS1" | u64 tag | u64 count | count x (i64 id | u32 len | bytes)
// Order of a save: values under a NEW tag, then the library's keys file carrying that tag (its rename is the commit
// point), then older values files are removed.  A kill at any moment leaves a pair that belongs together.
// Caller holds s.mu.
func (s *Store) save(path string) error {
	var tb [8]byte
	if _, err := rand.Read(tb[:]); err != nil {
		return err
	}
	tag := binary.LittleEndian.Uint64(tb[:]) | 1
	vpath := fmt.Sprintf("%s.values.%016x", path, tag)
	f, err := os.Create(vpath + ".tmp")
	if err != nil {
		return err
	}
	w := bufio.NewWriter(f)
	w.WriteString("TFPVALS1")
	binary.Write(w, binary.LittleEndian, tag)
	binary.Write(w, binary.LittleEndian, uint64(len(s.values)))
	for id, v := range s.values {
		binary.Write(w, binary.LittleEndian, id)
		binary.Write(w, binary.LittleEndian, uint32(len(v)))
		w.Write(v)
	}
	if err := w.Flush(); err != nil {
		f.Close()
		return err
	}
	if err := f.Sync(); err != nil {
		f.Close()
		return err
	}
	f.Close()
	if err := os.Rename(vpath+".tmp", vpath); err != nil {
		return err
	}
	cpath := C.CString(path)
	defer C.free(unsafe.Pointer(cpath))
	if err := call(func() C.int { return C.tfp_store_save_tagged(s.h, cpath, C.uint64_t(tag)) }); err != nil {
		return err
	}
	old, _ := filepath.Glob(path + ".values.*")
	for _, o := range old {
		if o != vpath {
			os.Remove(o)
		}
	}
	return nil
}

// Caller holds s.mu; the store must be empty.
func (s *Store) restore(path string) error {
	cpath := C.CString(path)
	defer C.free(unsafe.Pointer(cpath))
	var tag C.uint64_t
	if err := call(func() C.int { return C.tfp_snapshot_tag(cpath, &tag) }); err != nil {
		return err
	}
	f, err := os.Open(fmt.Sprintf("%s.values.%016x", path, uint64(tag)))
	if err != nil {
		return err
	}
	defer f.Close()
	r := bufio.NewReader(f)
	head := make([]byte, 24)
	if _, err := io.ReadFull(r, head); err != nil || string(head[:8]) != "TFPVALS1" {
		return errors.New("not a values file")
	}
	if binary.LittleEndian.Uint64(head[8:16]) != uint64(tag) {
		return errors.New("the values file does not belong to this snapshot")
	}
	count := binary.LittleEndian.Uint64(head[16:24])
	values := make(map[int64][]byte, count)
	rec := make([]byte, 12)
	for i := uint64(0); i < count; i++ {
		if _, err := io.ReadFull(r, rec); err != nil {
			return errors.New("values file truncated")
		}
		id := int64(binary.LittleEndian.Uint64(rec[:8]))
		v := make([]byte, binary.LittleEndian.Uint32(rec[8:12]))
		if _, err := io.ReadFull(r, v); err != nil {
			return errors.New("values file truncated")
		}
		values[id] = v
	}
	if err := call(func() C.int { return C.tfp_store_load(s.h, cpath) }); err != nil {
		return err
	}
	s.values = values
	return nil
}
