// core_cgo.go -- the cgo shim a maintainer of cjg/aliyun-gpushare-device-plugin would drop into
// pkg/gpu/nvidia/ to put libgpushare_core.so under the existing Go plugin.
//
// UNVERIFIED: neither the build image nor the GPU image has a Go toolchain, so this file has never
// been compiled. It is kept in sync with INTEGRATION.md (same code) and with include/gpushare.h by
// hand. What IS exercised is the same C ABI through ctypes and the C++ face (tests/).
//
// The rest of the integration (raw-bytes gRPC codec, Allocate keeping m.Lock() and the PATCH with
// its one retry) is described in INTEGRATION.md.
package nvidia

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../lib -lgpushare_core -ldl -lpthread
#include <stdlib.h>
#include "gpushare.h"
*/
import "C"

import (
	"fmt"
	"runtime"
	"unsafe"

	v1 "k8s.io/api/core/v1"
)

type coreTable struct{ t *C.gs_table }

// getDevices (nvidia.go:50-86): enumerate through NVML, build the id table once.
func buildTable(unit MemoryUnit) (*coreTable, error) {
	if rc := C.gs_nvml_open(nil); rc != C.GS_OK {
		return nil, fmt.Errorf("nvml: %s", C.GoString(C.gs_nvml_last_error()))
	}
	var n C.uint32_t
	C.gs_nvml_device_count(&n)
	gpus := make([]C.gs_gpu_info, int(n))
	for i := range gpus {
		if rc := C.gs_nvml_device_info(C.uint32_t(i), &gpus[i]); rc != C.GS_OK {
			return nil, fmt.Errorf("nvml: %s", C.GoString(C.gs_nvml_last_error()))
		}
	}
	u := C.int(C.GS_UNIT_GIB)
	if unit == MiBPrefix {
		u = C.GS_UNIT_MIB
	}
	var t *C.gs_table
	if rc := C.gs_table_build_from_gpus(&gpus[0], n, u, &t); rc != C.GS_OK {
		return nil, fmt.Errorf("gs_table_build_from_gpus: %s", C.GoString(C.gs_strerror(rc)))
	}
	return &coreTable{t}, nil
}

// ListAndWatch (server.go:180-193): borrowed, pre-encoded segments; no per-device work.
// The stream is registered with a raw codec (grpc.ForceServerCodec) whose Marshal returns
// the []byte it is given.
func (c *coreTable) listAndWatchBytes() []byte {
	gpus := int(C.gs_table_gpu_count(c.t))
	if gpus == 0 {
		return nil // a node without GPUs advertises an empty list (an empty message)
	}
	iov := make([]C.gs_iovec, gpus)
	var n C.uint32_t
	C.gs_law_payload(c.t, &iov[0], C.uint32_t(len(iov)), &n)
	out := make([]byte, 0, int(C.gs_law_payload_size(c.t)))
	for _, s := range iov[:n] {
		out = append(out, unsafe.Slice((*byte)(unsafe.Pointer(s.base)), int(s.len))...)
	}
	return out
}

// gsStr points a gs_str at a Go string's bytes. The views array is Go memory that holds these
// pointers, so each one is pinned (runtime.Pinner, Go 1.21) for the duration of the call.
func gsStr(pin *runtime.Pinner, s string) C.gs_str {
	if len(s) == 0 {
		return C.gs_str{p: (*C.char)(unsafe.Pointer(&emptyByte)), n: 0} // present but empty
	}
	d := unsafe.StringData(s)
	pin.Pin(d)
	return C.gs_str{p: (*C.char)(unsafe.Pointer(d)), n: C.size_t(len(s))}
}

var emptyByte byte

// Allocate (allocate.go:43-201). req is the request exactly as it came off the wire
// (raw codec); pods is what getPendingPodsInNode's lister returned, unfiltered.
func (m *NvidiaDevicePlugin) allocateCore(req []byte, pods []v1.Pod, nowUnixNano int64) (resp, patch []byte, res C.gs_alloc_result, err error) {
	var pin runtime.Pinner
	defer pin.Unpin()
	views := make([]C.gs_pod_view, len(pods))
	for i := range pods {
		p := &pods[i]
		views[i].uid, views[i].ns = gsStr(&pin, string(p.UID)), gsStr(&pin, p.Namespace)
		views[i].name, views[i].node_name = gsStr(&pin, p.Name), gsStr(&pin, p.Spec.NodeName)
		views[i].gpu_mem_limit = C.uint64_t(getGPUMemoryFromPodResource(p)) // podutils.go:122-131
		if v, ok := p.Annotations[EnvResourceIndex]; ok {
			views[i].ann_idx = gsStr(&pin, v)
		}
		if v, ok := p.Annotations[EnvResourceAssumeTime]; ok {
			views[i].ann_assume_time = gsStr(&pin, v)
		}
		if v, ok := p.Annotations[EnvAssignedFlag]; ok {
			views[i].ann_assigned = gsStr(&pin, v)
		}
	}
	ctx := C.gs_alloc_ctx{table: m.core.t, node_name: gsStr(&pin, nodeName), now_unix_nano: C.int64_t(nowUnixNano)}
	if m.disableCGPUIsolation {
		ctx.disable_cgpu_isolation = 1
	}
	resp = make([]byte, 512+256*8)
	patch = make([]byte, 192)
	var vp *C.gs_pod_view
	if len(views) > 0 {
		vp = &views[0]
	}
	call := func() C.int {
		return C.gs_allocate(&ctx, (*C.uint8_t)(unsafe.SliceData(req)), C.size_t(len(req)), vp, C.size_t(len(views)),
			(*C.uint8_t)(&resp[0]), C.size_t(len(resp)), (*C.char)(unsafe.Pointer(&patch[0])), C.size_t(len(patch)), &res)
	}
	rc := call()
	if rc == C.GS_E_NOSPACE {
		// a pod with many containers (about 230 bytes of response each): the required sizes came
		// back in res; gs_allocate is stateless, so calling again decides the same thing
		if int(res.resp_len) > len(resp) {
			resp = make([]byte, int(res.resp_len))
		}
		if int(res.patch_len) > len(patch) {
			patch = make([]byte, int(res.patch_len))
		}
		rc = call()
	}
	if rc != C.GS_OK {
		return nil, nil, res, fmt.Errorf("gs_allocate: %s", C.GoString(C.gs_strerror(rc)))
	}
	return resp[:res.resp_len], patch[:res.patch_len], res, nil
}
