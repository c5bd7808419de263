package main

// GPU-backed replacement for backend/go/stores/store.go.
//
// Same type name, same embedded mixin idea, same methods and error texts as the
// reference Store; every loop over keys happens behind cgo in libtfpstore.so
// (include/tfp_store.h).  Values stay in Go: the library hands out int64 ids and
// this file keeps id -> []byte.
//
// NOT COMPILED in the build image (no Go toolchain, no generated pb.go): the
// Python host localai_tfp_b200/store.py is the tested mirror of this file.

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../localai_tfp_b200 -ltfpstore
#include <stdlib.h>
#include "tfp_store.h"
*/
import "C"

import (
	"fmt"
	"sync"
	"unsafe"

	"github.com/mudler/LocalAI/pkg/grpc/base"
	pb "github.com/mudler/LocalAI/pkg/grpc/proto"
)

type Store struct {
	// base.Base (Locking() == false) instead of the reference's base.SingleThread:
	// the library serialises mutations itself and lets finds overlap.
	base.Base

	mu     sync.RWMutex
	h      *C.tfp_store
	values map[int64][]byte
}

func NewStore() *Store {
	s := &Store{values: map[int64][]byte{}}
	cfg := C.tfp_cfg{device: 0}
	if rc := C.tfp_store_open(&cfg, &s.h); rc != 0 {
		panic("tfp_store_open: " + C.GoString(C.tfp_last_error()))
	}
	return s
}

func lastErr() error { return fmt.Errorf("%s", C.GoString(C.tfp_last_error())) }

// pack copies the request's keys into one C block (cgo forbids lending Go
// pointers to Go pointers; the library copies again into pinned staging).
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
	buf := (*C.float)(C.malloc(C.size_t(n*dim*4 + 4)))
	dst := unsafe.Slice((*float32)(unsafe.Pointer(buf)), n*dim)
	for i, k := range keys {
		copy(dst[i*dim:(i+1)*dim], k.Floats)
	}
	return buf, dim, func() { C.free(unsafe.Pointer(buf)) }, nil
}

func (s *Store) Load(opts *pb.ModelOptions) error { return nil }

func (s *Store) StoresSet(opts *pb.StoresSetOptions) error {
	buf, dim, free, err := pack(opts.Keys)
	defer free()
	if err != nil {
		return err
	}
	n := len(opts.Keys)
	ids := make([]int64, n+1)
	s.mu.Lock()
	defer s.mu.Unlock()
	rc := C.tfp_store_set(s.h, buf, C.int64_t(n), C.int(dim), C.int64_t(len(opts.Values)),
		(*C.int64_t)(unsafe.Pointer(&ids[0])), nil)
	if rc != 0 {
		return lastErr()
	}
	for i := 0; i < n; i++ {
		s.values[ids[i]] = opts.Values[i].Bytes
	}
	return nil
}

func (s *Store) StoresDelete(opts *pb.StoresDeleteOptions) error {
	buf, dim, free, err := pack(opts.Keys)
	defer free()
	if err != nil {
		return err
	}
	n := len(opts.Keys)
	ids := make([]int64, n+1)
	s.mu.Lock()
	defer s.mu.Unlock()
	if rc := C.tfp_store_delete(s.h, buf, C.int64_t(n), C.int(dim), (*C.int64_t)(unsafe.Pointer(&ids[0]))); rc != 0 {
		return lastErr()
	}
	for i := 0; i < n; i++ {
		if ids[i] >= 0 {
			delete(s.values, ids[i])
		}
	}
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
	s.mu.RLock()
	defer s.mu.RUnlock()
	rc := C.tfp_store_get(s.h, buf, C.int64_t(n), C.int(dim), (*C.int64_t)(unsafe.Pointer(&ids[0])),
		(*C.int64_t)(unsafe.Pointer(&order[0])), &found)
	if rc != 0 {
		return pb.StoresGetResult{}, lastErr()
	}
	res := pb.StoresGetResult{}
	for j := 0; j < int(found); j++ {
		i := order[j]
		res.Keys = append(res.Keys, &pb.StoresKey{Floats: opts.Keys[i].Floats})
		res.Values = append(res.Values, &pb.StoresValue{Bytes: s.values[ids[i]]})
	}
	return res, nil
}

func (s *Store) StoresFind(opts *pb.StoresFindOptions) (pb.StoresFindResult, error) {
	q := opts.Key.Floats
	dim := len(q)
	topk := int(opts.TopK)
	s.mu.RLock()
	defer s.mu.RUnlock()
	var st C.tfp_stats
	C.tfp_store_stats(s.h, &st)
	slots := topk
	if int64(slots) > int64(st.rows) {
		slots = int(st.rows) // the result holds min(TopK, rows) entries
	}
	if slots < 1 {
		slots = 1
	}
	call := topk
	if topk >= 1 {
		call = slots
	}
	ids := make([]int64, slots)
	sims := make([]float32, slots)
	keys := make([]float32, slots*dim+1)
	var n C.int
	var qp *C.float
	if dim > 0 {
		qp = (*C.float)(unsafe.Pointer(&q[0]))
	}
	rc := C.tfp_store_find(s.h, qp, C.int(dim), C.int(call), &n, (*C.int64_t)(unsafe.Pointer(&ids[0])),
		(*C.float)(unsafe.Pointer(&sims[0])), (*C.float)(unsafe.Pointer(&keys[0])), nil)
	if rc != 0 {
		return pb.StoresFindResult{}, lastErr()
	}
	res := pb.StoresFindResult{Similarities: sims[:int(n)]}
	for i := 0; i < int(n); i++ {
		res.Keys = append(res.Keys, &pb.StoresKey{Floats: keys[i*dim : (i+1)*dim]})
		res.Values = append(res.Values, &pb.StoresValue{Bytes: s.values[ids[i]]})
	}
	return res, nil
}
