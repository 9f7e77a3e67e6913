// Package store — GPU provider for the reference's Store interface.
//
// NOT COMPILED IN THIS REPOSITORY'S ENVIRONMENT (no Go toolchain in the image).  It is the
// binding a doc-agents maintainer drops into internal/store/gpu.go next to postgres.go (same
// package, so it reaches PostgresStore.db); integration/deps_gpu.patch and
// integration/config_gpu.patch add the provider case and its env knobs.  It is kept mechanical
// on purpose: type conversion and one C call per method.  Every behaviour it relies on below the
// C boundary is verified by tests/ (ctypes twin: doc_agents_b200/_lib.py, store.py; C++ twin:
// include/docagents_b200.hpp + tests/cpp/contract_test.cpp, which mirror this file line by line).
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
	"runtime"
	"strconv"
	"strings"
	"sync"
	"unsafe"

	"github.com/google/uuid"
	"github.com/lib/pq"

	"doc-agents/internal/embeddings"
)

// GPUStore decorates the Postgres store: relational methods are delegated unchanged,
// SaveChunks is observed (chunk -> document), SaveEmbeddings writes through to Postgres and
// to the device, TopK searches on the device and hydrates <= k rows from Postgres.
//
// Postgres stays the system of record.  The device store is a cache of `embeddings JOIN chunks`:
// it is filled at start (Preload), by this process's own SaveEmbeddings, and -- because the
// writer (analysis agent) and the reader (query agent) are different processes
// (docker-compose.yml:87-128) -- on demand: TopK first asks the device how many rows it holds for
// each requested document (da_store_doc_rows) and loads the documents it has never seen.
type GPUStore struct {
	*PostgresStore
	h        *C.da_store
	dim      int
	readOnly bool // attached to another process's store: saves go to Postgres only, the writer's process feeds the device

	mu    sync.Mutex
	stale map[uuid.UUID]struct{} // documents whose device copy may be behind Postgres (a device upsert failed)
}

type GPUOptions struct {
	Dim           int     // GPU_DIM, default 3072 (vector(3072): postgres.go:84)
	Devices       []int   // GPU_DEVICES
	CapacityRows  uint64  // GPU_CAPACITY_ROWS, per device
	MinSimilarity float64 // GPU_MIN_SIMILARITY, default 0.7: a float64 like the reference's constant (postgres.go:223), compared as (double)score >= floor
	Int8Sidecar   bool    // GPU_INT8_SIDECAR: keep an int8 copy of every row (25 % more HBM, ~3.6x queries/s, same answers)
	BatchWindowUs uint32  // GPU_BATCH_WINDOW_US: merge TopK calls that arrive within this window into one pass (0 = off)
	Preload       bool    // GPU_PRELOAD: stream embeddings JOIN chunks into the device store at start
	// One device store for both agents (same GPU, two processes): the analysis agent creates the store and exports
	// it under SharedName (da_store_export); the query agent sets Attach and maps the same HBM read-only
	// (da_store_attach) -- every TopK there first replays what the writer has saved since.
	SharedName string // GPU_SHARED_NAME, e.g. "/docagents-gpu0"; empty = a private store
	Attach     bool   // GPU_ATTACH: this process only reads the store exported under SharedName
}

// call runs one C entry point and, if it failed, fetches its message.  da_last_error() is
// thread-local in the library, and a goroutine may be moved to another OS thread between two cgo
// calls: the thread is pinned for the pair.
func call(f func() C.int) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	rc := f()
	if rc == C.DA_OK {
		return nil
	}
	return fmt.Errorf("gpu store: %s (code %d)", C.GoString(C.da_last_error()), int(rc))
}

// NewGPU is what buildStore calls for STORE_PROVIDER=gpu (internal/app/deps.go:179-202).
func NewGPU(ctx context.Context, pg *PostgresStore, o GPUOptions) (*GPUStore, error) {
	if len(o.Devices) == 0 {
		o.Devices = []int{0}
	}
	var cfg C.da_config
	cfg.struct_size = C.uint32_t(unsafe.Sizeof(cfg))
	cfg.dim = C.int32_t(o.Dim)
	cfg.n_devices = C.int32_t(len(o.Devices))
	for i, d := range o.Devices {
		cfg.device_ids[i] = C.int32_t(d)
	}
	cfg.capacity_rows = C.uint64_t(o.CapacityRows)
	cfg.min_similarity = C.float(o.MinSimilarity)
	cfg.min_similarity_f64 = C.double(o.MinSimilarity)
	if o.Int8Sidecar {
		cfg.flags |= C.DA_STORE_Q8_SIDECAR
	}
	cfg.batch_window_us = C.uint32_t(o.BatchWindowUs)
	s := &GPUStore{PostgresStore: pg, dim: o.Dim, stale: map[uuid.UUID]struct{}{}, readOnly: o.Attach}
	if o.Attach {
		name := C.CString(o.SharedName)
		defer C.free(unsafe.Pointer(name))
		if err := call(func() C.int { return C.da_store_attach(name, &s.h) }); err != nil {
			return nil, err
		}
		return s, nil // the exporting process loads and saves; this one follows its log
	}
	if err := call(func() C.int { return C.da_store_create(&cfg, &s.h) }); err != nil {
		return nil, err
	}
	if o.SharedName != "" {
		name := C.CString(o.SharedName)
		defer C.free(unsafe.Pointer(name))
		if err := call(func() C.int { return C.da_store_export(s.h, name) }); err != nil {
			s.Close()
			return nil, err
		}
	}
	if o.Preload {
		if err := s.load(ctx, nil); err != nil { // every document
			s.Close()
			return nil, fmt.Errorf("gpu store: cold start: %w", err)
		}
	}
	return s, nil
}

func (s *GPUStore) Close() { C.da_store_destroy(s.h) }

// load streams `embeddings JOIN chunks` (chunk id, document id, vector) for the given documents
// -- nil = all -- into da_store_bulk_load, a few thousand rows per call, ordered by document so
// that a document's rows arrive together.  Vectors cross the SQL boundary as pgvector text
// (database/sql has no binary vector type); parsing them back with ParseFloat(.., 32) restores the
// exact float32 bits that vectorToString wrote (postgres.go:321-330).  The upsert on the device is
// keyed by chunk id, so loading a row twice is harmless.
func (s *GPUStore) load(ctx context.Context, docIDs []uuid.UUID) error {
	q := `SELECT e.chunk_id, c.document_id, e.vector::text
	        FROM embeddings e JOIN chunks c ON c.id = e.chunk_id`
	args := []any{}
	if docIDs != nil {
		q += ` WHERE c.document_id = ANY($1)`
		args = append(args, pqUUIDArray(docIDs))
	}
	q += ` ORDER BY c.document_id, c.ord`
	rows, err := s.PostgresStore.db.QueryContext(ctx, q, args...)
	if err != nil {
		return err
	}
	defer rows.Close()
	const batchRows = 4096
	chunkIDs := make([]byte, 0, 16*batchRows)
	docs := make([]byte, 0, 16*batchRows)
	vecs := make([]float32, 0, s.dim*batchRows)
	flush := func() error {
		n := len(chunkIDs) / 16
		if n == 0 {
			return nil
		}
		err := call(func() C.int {
			return C.da_store_bulk_load(s.h, (*C.uint8_t)(unsafe.Pointer(&chunkIDs[0])),
				(*C.uint8_t)(unsafe.Pointer(&docs[0])), (*C.float)(unsafe.Pointer(&vecs[0])),
				C.size_t(n), C.int32_t(s.dim))
		})
		chunkIDs, docs, vecs = chunkIDs[:0], docs[:0], vecs[:0]
		return err
	}
	for rows.Next() {
		var cid, did uuid.UUID
		var text string
		if err := rows.Scan(&cid, &did, &text); err != nil {
			return err
		}
		fields := strings.Split(strings.Trim(text, "[]"), ",")
		if len(fields) != s.dim {
			return fmt.Errorf("embedding of chunk %s has %d dimensions, the store %d", cid, len(fields), s.dim)
		}
		for _, f := range fields {
			v, err := strconv.ParseFloat(f, 32)
			if err != nil {
				return err
			}
			vecs = append(vecs, float32(v))
		}
		chunkIDs = append(chunkIDs, cid[:]...)
		docs = append(docs, did[:]...)
		if len(chunkIDs) == 16*batchRows {
			if err := flush(); err != nil {
				return err
			}
		}
	}
	if err := rows.Err(); err != nil {
		return err
	}
	return flush()
}

// SaveChunks: Postgres assigns the chunk ids (postgres.go:150-157); the device store only
// needs chunk -> document.
func (s *GPUStore) SaveChunks(ctx context.Context, docID uuid.UUID, chunks []Chunk) ([]Chunk, error) {
	out, err := s.PostgresStore.SaveChunks(ctx, docID, chunks)
	if err != nil || len(out) == 0 || s.readOnly {
		return out, err
	}
	ids := make([]byte, 16*len(out))
	for i, c := range out {
		copy(ids[16*i:], c.ID[:])
	}
	err = call(func() C.int {
		return C.da_store_save_chunks(s.h, (*C.uint8_t)(unsafe.Pointer(&docID[0])),
			(*C.uint8_t)(unsafe.Pointer(&ids[0])), C.size_t(len(out)))
	})
	return out, err
}

// SaveEmbeddings: write-through, Postgres first (the system of record), the device second.  The
// C call copies everything it needs before returning (no Go pointer is retained), and replaces
// vectorToString's text encoding (postgres.go:321-330) with the raw float32 bits.
//
// Failure policy when Postgres succeeded and the device upsert did not: the error is returned
// (the analysis handler fails the task, and the queue retries it up to five times:
// internal/queue/nats.go:69-83; both upserts are idempotent -- ON CONFLICT (chunk_id) DO UPDATE on
// one side, upsert by chunk id on the other -- so a retry converges), and the documents touched
// are marked stale: until a later load succeeds, TopK for them is answered by PostgresStore.TopK,
// never from a device copy that may be behind.  A process in which this store never saw the
// chunks (SaveChunks ran in the parser's process) announces them from Postgres first.
func (s *GPUStore) SaveEmbeddings(ctx context.Context, embs []Embedding) error {
	if len(embs) == 0 {
		return nil // postgres.go:177-179
	}
	if err := s.PostgresStore.SaveEmbeddings(ctx, embs); err != nil {
		return err
	}
	if s.readOnly {
		return nil
	}
	dim := len(embs[0].Vector)
	ids := make([]byte, 16*len(embs))
	vecs := make([]float32, dim*len(embs))
	chunkIDs := make([]uuid.UUID, len(embs))
	for i, e := range embs {
		if len(e.Vector) != dim {
			return errors.New("gpu store: embeddings of different dimensions in one batch")
		}
		copy(ids[16*i:], e.ChunkID[:])
		copy(vecs[dim*i:], e.Vector)
		chunkIDs[i] = e.ChunkID
	}
	docOf, err := s.documentsOf(ctx, chunkIDs) // chunk -> document, from Postgres (the parser saved them)
	if err == nil {
		err = s.announce(docOf)
	}
	if err == nil {
		err = call(func() C.int {
			return C.da_store_save_embeddings(s.h, (*C.uint8_t)(unsafe.Pointer(&ids[0])),
				(*C.float)(unsafe.Pointer(&vecs[0])), C.size_t(len(embs)), C.int32_t(dim))
		})
	}
	if err != nil {
		s.mu.Lock()
		for _, d := range docOf {
			s.stale[d] = struct{}{}
		}
		s.mu.Unlock()
		return err
	}
	return nil
}

// documentsOf returns chunk id -> document id for the given chunks.
func (s *GPUStore) documentsOf(ctx context.Context, chunkIDs []uuid.UUID) (map[uuid.UUID]uuid.UUID, error) {
	rows, err := s.PostgresStore.db.QueryContext(ctx, `SELECT id, document_id FROM chunks WHERE id = ANY($1)`, pqUUIDArray(chunkIDs))
	if err != nil {
		return nil, err
	}
	defer rows.Close()
	out := make(map[uuid.UUID]uuid.UUID, len(chunkIDs))
	for rows.Next() {
		var c, d uuid.UUID
		if err := rows.Scan(&c, &d); err != nil {
			return nil, err
		}
		out[c] = d
	}
	return out, rows.Err()
}

// announce tells the device store which document each chunk belongs to (da_store_save_chunks per document).
func (s *GPUStore) announce(docOf map[uuid.UUID]uuid.UUID) error {
	byDoc := map[uuid.UUID][]byte{}
	for c, d := range docOf {
		byDoc[d] = append(byDoc[d], c[:]...)
	}
	for d, ids := range byDoc {
		d, ids := d, ids
		if err := call(func() C.int {
			return C.da_store_save_chunks(s.h, (*C.uint8_t)(unsafe.Pointer(&d[0])),
				(*C.uint8_t)(unsafe.Pointer(&ids[0])), C.size_t(len(ids)/16))
		}); err != nil {
			return err
		}
	}
	return nil
}

// ensureResident loads the requested documents this process has never been given (embedded by the
// analysis agent, another process) and re-loads stale ones.  Returns the documents that are still
// not trustworthy on the device (their load failed).
func (s *GPUStore) ensureResident(ctx context.Context, docIDs []uuid.UUID) (map[uuid.UUID]struct{}, error) {
	if len(docIDs) == 0 {
		return nil, nil
	}
	docs := make([]byte, 16*len(docIDs))
	for i, d := range docIDs {
		copy(docs[16*i:], d[:])
	}
	counts := make([]uint32, len(docIDs))
	if err := call(func() C.int {
		return C.da_store_doc_rows(s.h, (*C.uint8_t)(unsafe.Pointer(&docs[0])), C.size_t(len(docIDs)),
			(*C.uint32_t)(unsafe.Pointer(&counts[0])))
	}); err != nil {
		return nil, err
	}
	var want []uuid.UUID
	s.mu.Lock()
	for i, d := range docIDs {
		if _, st := s.stale[d]; st || counts[i] == 0 {
			want = append(want, d)
		}
	}
	s.mu.Unlock()
	if len(want) == 0 {
		return nil, nil
	}
	if s.readOnly { // only the exporting process can load: answer these from Postgres until it has them
		bad := make(map[uuid.UUID]struct{}, len(want))
		for _, d := range want {
			bad[d] = struct{}{}
		}
		return bad, nil
	}
	if err := s.load(ctx, want); err != nil {
		bad := make(map[uuid.UUID]struct{}, len(want))
		for _, d := range want {
			bad[d] = struct{}{}
		}
		return bad, nil // the caller answers from Postgres this time
	}
	s.mu.Lock()
	for _, d := range want {
		delete(s.stale, d)
	}
	s.mu.Unlock()
	return nil, nil
}

// TopK: device search, then one hydration query for <= k chunk ids, re-ordered by device rank.
func (s *GPUStore) TopK(ctx context.Context, docIDs []uuid.UUID, vector embeddings.Vector, k int) ([]SearchResult, error) {
	if len(vector) == 0 {
		// the reference sends '[]' to pgvector, which rejects it: an error, not an empty answer
		return nil, errors.New("gpu store: vector must have at least 1 dimension")
	}
	if len(docIDs) == 0 {
		return nil, nil // document_id = ANY('{}') matches nothing (postgres.go:239)
	}
	bad, err := s.ensureResident(ctx, docIDs)
	if err != nil {
		return nil, err
	}
	if len(bad) > 0 {
		return s.PostgresStore.TopK(ctx, docIDs, vector, k) // never answer from a device copy known to be behind
	}
	kk := k
	if kk < 1 {
		kk = 1
	}
	docs := make([]byte, 16*len(docIDs))
	for i, d := range docIDs {
		copy(docs[16*i:], d[:])
	}
	outIDs := make([]byte, 16*kk)
	outScores := make([]float32, kk)
	var n C.int32_t
	if err := call(func() C.int {
		return C.da_store_topk(s.h, (*C.uint8_t)(unsafe.Pointer(&docs[0])), C.size_t(len(docIDs)),
			(*C.float)(unsafe.Pointer(&vector[0])), C.int32_t(len(vector)), C.int32_t(k),
			(*C.uint8_t)(unsafe.Pointer(&outIDs[0])), (*C.float)(unsafe.Pointer(&outScores[0])), &n)
	}); err != nil {
		return nil, err // -> "search failed", HTTP 500 (cmd/query/main.go:102-105)
	}
	if n == 0 {
		return nil, nil // the reference returns a nil slice for zero rows (postgres.go:250,284)
	}
	ids := make([]uuid.UUID, int(n))
	for i := range ids {
		copy(ids[i][:], outIDs[16*i:16*i+16])
	}
	byID, err := s.hydrateChunks(ctx, ids)
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
	if len(results) == 0 {
		return nil, nil
	}
	return results, nil
}

// hydrateChunks fetches, for <= k chunk ids, the columns PostgresStore.TopK returns beside the
// score (postgres.go:226-236): the chunk, and its document's summary through the same LEFT JOIN.
func (s *GPUStore) hydrateChunks(ctx context.Context, ids []uuid.UUID) (map[uuid.UUID]SearchResult, error) {
	rows, err := s.PostgresStore.db.QueryContext(ctx, `
		SELECT c.id, c.document_id, c.ord, c.text, c.token_count,
		       COALESCE(s.summary, ''), COALESCE(s.key_points, ARRAY[]::TEXT[])
		FROM chunks c
		LEFT JOIN summaries s ON s.document_id = c.document_id
		WHERE c.id = ANY($1)`, pqUUIDArray(ids))
	if err != nil {
		return nil, err
	}
	defer rows.Close()
	out := make(map[uuid.UUID]SearchResult, len(ids))
	for rows.Next() {
		var (
			chunkID, docID uuid.UUID
			ord, tokens    int
			text, summary  string
			keyPoints      []string
		)
		if err := rows.Scan(&chunkID, &docID, &ord, &text, &tokens, &summary, pq.Array(&keyPoints)); err != nil {
			return nil, err
		}
		out[chunkID] = SearchResult{
			Chunk:   Chunk{ID: chunkID, DocumentID: docID, Index: ord, Text: text, TokenCount: tokens},
			Summary: Summary{DocumentID: docID, Summary: summary, KeyPoints: keyPoints},
		}
	}
	return out, rows.Err()
}
