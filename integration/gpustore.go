// Package store — GPU provider for the reference's Store interface.
//
// NOT COMPILED IN THIS REPOSITORY'S ENVIRONMENT (no Go toolchain in the image).  It is the
// binding a doc-agents maintainer drops into internal/store/ next to postgres.go; it is kept
// mechanical on purpose: type conversion and one C call per method.  The behaviour it relies
// on is verified at the C ABI by tests/ (ctypes twin: doc_agents_b200/_lib.py, store.py).
//
// Build: CGO_ENABLED=1 (the reference Dockerfile sets 0 at Dockerfile:10), a CUDA runtime base
// image instead of distroless (Dockerfile:12), and libdocagents_b200.so + libnccl on the
// loader path.
package store

/*
#cgo CFLAGS: -I${SRCDIR}/../../third_party/docagents_b200/include
#cgo LDFLAGS: -L${SRCDIR}/../../third_party/docagents_b200/lib -ldocagents_b200
#include <stdlib.h>
#include "docagents_b200.h"
*/
import "C"

import (
	"context"
	"errors"
	"fmt"
	"unsafe"

	"github.com/google/uuid"

	"doc-agents/internal/embeddings"
)

// GPUStore decorates the Postgres store: relational methods are delegated unchanged,
// SaveChunks is observed (chunk -> document), SaveEmbeddings writes through to Postgres and
// to the device, TopK searches on the device and hydrates <= k rows from Postgres.
type GPUStore struct {
	*PostgresStore
	h *C.da_store
}

type GPUOptions struct {
	Dim           int     // GPU_DIM, default 3072 (vector(3072): postgres.go:84)
	Devices       []int   // GPU_DEVICES
	CapacityRows  uint64  // GPU_CAPACITY_ROWS, per device
	MinSimilarity float32 // GPU_MIN_SIMILARITY, default 0.7 (postgres.go:223)
	Int8Sidecar   bool    // GPU_INT8_SIDECAR: keep an int8 copy of every row (25 % more HBM, ~3.6x queries/s, same answers)
	BatchWindowUs uint32  // GPU_BATCH_WINDOW_US: merge TopK calls that arrive within this window into one pass (0 = off)
}

func gpuErr(rc C.int) error {
	if rc == C.DA_OK {
		return nil
	}
	return fmt.Errorf("gpu store: %s (code %d)", C.GoString(C.da_last_error()), int(rc))
}

// NewGPU is what buildStore calls for STORE_PROVIDER=gpu (internal/app/deps.go:179-202).
func NewGPU(pg *PostgresStore, o GPUOptions) (*GPUStore, error) {
	var cfg C.da_config
	cfg.struct_size = C.uint32_t(unsafe.Sizeof(cfg))
	cfg.dim = C.int32_t(o.Dim)
	cfg.n_devices = C.int32_t(len(o.Devices))
	for i, d := range o.Devices {
		cfg.device_ids[i] = C.int32_t(d)
	}
	cfg.capacity_rows = C.uint64_t(o.CapacityRows)
	cfg.min_similarity = C.float(o.MinSimilarity)
	if o.Int8Sidecar {
		cfg.flags |= C.DA_STORE_Q8_SIDECAR
	}
	cfg.batch_window_us = C.uint32_t(o.BatchWindowUs)
	s := &GPUStore{PostgresStore: pg}
	if err := gpuErr(C.da_store_create(&cfg, &s.h)); err != nil {
		return nil, err
	}
	return s, nil
}

func (s *GPUStore) Close() { C.da_store_destroy(s.h) }

// SaveChunks: Postgres assigns the chunk ids (postgres.go:150-157); the device store only
// needs chunk -> document.
func (s *GPUStore) SaveChunks(ctx context.Context, docID uuid.UUID, chunks []Chunk) ([]Chunk, error) {
	out, err := s.PostgresStore.SaveChunks(ctx, docID, chunks)
	if err != nil || len(out) == 0 {
		return out, err
	}
	ids := make([]byte, 16*len(out))
	for i, c := range out {
		copy(ids[16*i:], c.ID[:])
	}
	rc := C.da_store_save_chunks(s.h, (*C.uint8_t)(unsafe.Pointer(&docID[0])),
		(*C.uint8_t)(unsafe.Pointer(&ids[0])), C.size_t(len(out)))
	return out, gpuErr(rc)
}

// SaveEmbeddings: write-through.  The C call copies everything it needs before returning
// (no Go pointer is retained), and replaces vectorToString's text encoding (postgres.go:321-330)
// with the raw float32 bits.
func (s *GPUStore) SaveEmbeddings(ctx context.Context, embs []Embedding) error {
	if len(embs) == 0 {
		return nil
	}
	if err := s.PostgresStore.SaveEmbeddings(ctx, embs); err != nil {
		return err
	}
	dim := len(embs[0].Vector)
	ids := make([]byte, 16*len(embs))
	vecs := make([]float32, dim*len(embs))
	for i, e := range embs {
		if len(e.Vector) != dim {
			return errors.New("gpu store: embeddings of different dimensions in one batch")
		}
		copy(ids[16*i:], e.ChunkID[:])
		copy(vecs[dim*i:], e.Vector)
	}
	return gpuErr(C.da_store_save_embeddings(s.h, (*C.uint8_t)(unsafe.Pointer(&ids[0])),
		(*C.float)(unsafe.Pointer(&vecs[0])), C.size_t(len(embs)), C.int32_t(dim)))
}

// TopK: device search, then one hydration query for <= k chunk ids, re-ordered by device rank.
func (s *GPUStore) TopK(ctx context.Context, docIDs []uuid.UUID, vector embeddings.Vector, k int) ([]SearchResult, error) {
	if k <= 0 || len(vector) == 0 {
		return nil, nil
	}
	var docPtr *C.uint8_t
	docs := make([]byte, 16*len(docIDs))
	for i, d := range docIDs {
		copy(docs[16*i:], d[:])
	}
	if len(docs) > 0 {
		docPtr = (*C.uint8_t)(unsafe.Pointer(&docs[0]))
	}
	outIDs := make([]byte, 16*k)
	outScores := make([]float32, k)
	var n C.int32_t
	rc := C.da_store_topk(s.h, docPtr, C.size_t(len(docIDs)), (*C.float)(unsafe.Pointer(&vector[0])),
		C.int32_t(len(vector)), C.int32_t(k), (*C.uint8_t)(unsafe.Pointer(&outIDs[0])),
		(*C.float)(unsafe.Pointer(&outScores[0])), &n)
	if err := gpuErr(rc); err != nil {
		return nil, err
	}
	if n == 0 {
		return nil, nil // the reference returns a nil slice for zero rows (postgres.go:250,284)
	}
	ids := make([]uuid.UUID, int(n))
	for i := range ids {
		copy(ids[i][:], outIDs[16*i:16*i+16])
	}
	// hydrate: SELECT c.id, c.document_id, c.ord, c.text, c.token_count, COALESCE(s.summary,''),
	// COALESCE(s.key_points, ARRAY[]::TEXT[]) FROM chunks c LEFT JOIN summaries s ON
	// s.document_id = c.document_id WHERE c.id = ANY($1)      -- the columns of postgres.go:226-236
	byID, err := s.PostgresStore.hydrateChunks(ctx, ids)
	if err != nil {
		return nil, err
	}
	results := make([]SearchResult, 0, len(ids))
	for i, id := range ids {
		r, ok := byID[id]
		if !ok {
			continue // chunk row deleted since the embedding was stored (inner join: postgres.go:237)
		}
		r.Score = outScores[i]
		results = append(results, r)
	}
	return results, nil
}
