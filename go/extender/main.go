// Scheduler-extender shim over libgputopo.so: the two HTTP verbs of the reference's policy file
// (design.md:92-113: urlPrefix, prioritizeVerb "sort", bindVerb "bind", nodeCacheCapable, managed resource
// aliyun.com/gpu) answered by the C-ABI of include/gputopo.h.
//
// NOT COMPILED IN THIS REPOSITORY'S IMAGE (no Go toolchain): shipped as source, as SURVEY.md section 8b asks.
// Everything with arithmetic in it lives behind the cgo calls; this file is transport only.  The node watch
// (annotations -> gt_load_nodes, design.md:78-82) and the pod patch after bind (design.md:223-234) need a
// Kubernetes client and are left as the two marked hooks.
//
//   CGO_CFLAGS=-I$REPO/include CGO_LDFLAGS="-L$REPO/gpu-topology-on-k8s_b200 -lgputopo" go build ./go/extender
package main

/*
#include <stdlib.h>
#include "gputopo.h"
*/
import "C"

import (
	"bytes"
	"fmt"
	"io"
	"log"
	"net/http"
	"strings"
	"sync"
	"time"
	"unsafe"
)

type allocator struct {
	mu    sync.Mutex // one caller at a time (gputopo.h); kube-scheduler sends one pod at a time anyway
	ctx   *C.gt_ctx
	nodes int
}

func newAllocator(device int, maxNodes uint32) (*allocator, error) {
	cfg := C.gt_config{struct_size: C.uint32_t(unsafe.Sizeof(C.gt_config{})), device: C.int32_t(device),
		max_nodes: C.uint32_t(maxNodes)}
	a := &allocator{}
	if rc := C.gt_create(&cfg, &a.ctx); rc != 0 {
		return nil, fmt.Errorf("gt_create: %s", C.GoString(C.gt_strerror(rc)))
	}
	return a, nil
}

func (a *allocator) err(rc C.int, what string) error {
	if rc == 0 {
		return nil
	}
	return fmt.Errorf("%s: %s: %s", what, C.GoString(C.gt_strerror(rc)), C.GoString(C.gt_last_error(a.ctx)))
}

// loadNodes: hook for the node informer.  annotations[i] is the text of node i's GPU_<ABBR>_<a>_<b> annotations
// (design.md:78-82), free[i] the complement of the ALIYUN_COM_GPU_GROUP ids of the pods running there (design.md:231).
func (a *allocator) loadNodes(names []string, annotations []string, free []uint16) error {
	a.mu.Lock()
	defer a.mu.Unlock()
	n := len(names)
	levels := make([]byte, 256*n)
	nGpus := make([]byte, n)
	for i, text := range annotations {
		cs := C.CString(text)
		var cnt C.uint32_t
		rc := C.gt_annotation_decode(cs, C.size_t(len(text)), (*C.uint8_t)(unsafe.Pointer(&levels[256*i])), &cnt)
		C.free(unsafe.Pointer(cs))
		if rc != 0 {
			return a.err(rc, "gt_annotation_decode "+names[i])
		}
		nGpus[i] = byte(cnt)
	}
	if rc := C.gt_load_nodes(a.ctx, 0, C.uint32_t(n), (*C.uint8_t)(unsafe.Pointer(&levels[0])),
		(*C.uint8_t)(unsafe.Pointer(&nGpus[0])), nil, nil, (*C.uint16_t)(unsafe.Pointer(&free[0]))); rc != 0 {
		return a.err(rc, "gt_load_nodes")
	}
	joined := strings.Join(names, "\n")
	cs := C.CString(joined)
	defer C.free(unsafe.Pointer(cs))
	a.nodes = n
	return a.err(C.gt_set_node_names(a.ctx, cs, C.size_t(len(joined))), "gt_set_node_names")
}

func (a *allocator) serveSort(w http.ResponseWriter, r *http.Request) {
	body, err := io.ReadAll(r.Body)
	if err != nil || len(body) == 0 {
		http.Error(w, "empty body", http.StatusBadRequest)
		return
	}
	out := make([]byte, 96*a.nodes+len(body)+64)
	var n C.size_t
	a.mu.Lock()
	rc := C.gt_ext_sort_json(a.ctx, (*C.char)(unsafe.Pointer(&body[0])), C.size_t(len(body)),
		(*C.char)(unsafe.Pointer(&out[0])), C.size_t(len(out)), &n)
	e := a.err(rc, "sort")
	a.mu.Unlock()
	if e != nil {
		http.Error(w, e.Error(), http.StatusBadRequest)
		return
	}
	w.Header().Set("Content-Type", "application/json")
	w.Write(out[:n])
}

func (a *allocator) serveBind(w http.ResponseWriter, r *http.Request) {
	body, err := io.ReadAll(r.Body)
	if err != nil || len(body) == 0 {
		http.Error(w, "empty body", http.StatusBadRequest)
		return
	}
	out, ann := make([]byte, 1024), make([]byte, 256)
	a.mu.Lock()
	rc := C.gt_ext_bind_json(a.ctx, (*C.char)(unsafe.Pointer(&body[0])), C.size_t(len(body)), C.int64_t(time.Now().Unix()),
		(*C.char)(unsafe.Pointer(&out[0])), C.size_t(len(out)), (*C.char)(unsafe.Pointer(&ann[0])), C.size_t(len(ann)))
	e := a.err(rc, "bind")
	a.mu.Unlock()
	if e != nil {
		http.Error(w, e.Error(), http.StatusBadRequest)
		return
	}
	if i := bytes.IndexByte(ann, 0); i > 0 {
		// hook: patch these "KEY: value" lines onto the pod, then create the Binding (design.md:223-234)
		log.Printf("bind annotations:\n%s", ann[:i])
	}
	w.Header().Set("Content-Type", "application/json")
	w.Write(out[:bytes.IndexByte(out, 0)])
}

// podDeleted: hook for the pod informer (a pod carrying ALIYUN_COM_GPU_GROUP went away)
func (a *allocator) podDeleted(uid string) error {
	a.mu.Lock()
	defer a.mu.Unlock()
	cs := C.CString(uid)
	defer C.free(unsafe.Pointer(cs))
	return a.err(C.gt_ext_pod_deleted(a.ctx, cs), "gt_ext_pod_deleted")
}

func main() {
	a, err := newAllocator(0, 100000)
	if err != nil {
		log.Fatal(err) // no CUDA device: there is no CPU fallback
	}
	defer C.gt_destroy(a.ctx)
	const prefix = "/gputopology-scheduler" // urlPrefix of design.md:98: http://127.0.0.1:32743/gputopology-scheduler
	http.HandleFunc(prefix+"/sort", a.serveSort)
	http.HandleFunc(prefix+"/bind", a.serveBind)
	log.Fatal(http.ListenAndServe("127.0.0.1:32743", nil))
}
