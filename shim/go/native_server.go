// native_server.go -- Start/Stop/Serve/Register of pkg/gpu/nvidia/server.go:114-177,232-249 on the
// library's own gRPC transport (gs_server_*, gs_register) instead of grpc-go: the Go side keeps the
// pod lister and the PATCH (client-go) and is called back for each Allocate.
//
// UNVERIFIED: no Go toolchain on either image (see core_cgo.go). The same three C calls are
// exercised from C++ (tests/cpp/test_plugin.cpp, daemon_cpp/main.cpp) and Python.
package nvidia

/*
#include <stdlib.h>
#include "gpushare.h"

// cgo cannot take the address of a Go function: this trampoline is the gs_server_allocate_fn.
extern int goAllocate(void *user, uint8_t *req, size_t req_len, uint8_t *resp, size_t cap, size_t *resp_len);
static int allocate_trampoline(void *user, const uint8_t *req, size_t req_len, uint8_t *resp,
                               size_t cap, size_t *resp_len) {
  return goAllocate(user, (uint8_t *)req, req_len, resp, cap, resp_len);
}
static gs_server_allocate_fn allocate_fn(void) { return allocate_trampoline; }
*/
import "C"

import (
	"fmt"
	"runtime/cgo"
	"unsafe"
)

type nativeServer struct {
	srv    *C.gs_server
	handle cgo.Handle
	sock   *C.char
}

// allocateBody is what allocate.go:43-201 becomes: list pods, C.gs_allocate, PATCH (see core_cgo.go).
type allocateBody func(req []byte, resp []byte) (n int, status C.int)

//export goAllocate
func goAllocate(user unsafe.Pointer, req *C.uint8_t, reqLen C.size_t, resp *C.uint8_t, cap C.size_t,
	respLen *C.size_t) C.int {
	body := cgo.Handle(user).Value().(allocateBody)
	n, st := body(unsafe.Slice((*byte)(unsafe.Pointer(req)), int(reqLen)),
		unsafe.Slice((*byte)(unsafe.Pointer(resp)), int(cap)))
	*respLen = C.size_t(n)
	return st
}

// Serve (server.go:232-249): Start + Register; a failed Register stops the server again.
func serveNative(plugin *C.gs_plugin, socket, kubeletSocket, resource string, body allocateBody) (*nativeServer, error) {
	ns := &nativeServer{handle: cgo.NewHandle(body), sock: C.CString(socket)}
	var cfg C.gs_server_config
	cfg.socket_path = ns.sock
	cfg.allocate = C.allocate_fn()
	cfg.user = unsafe.Pointer(ns.handle)
	if rc := C.gs_server_start(plugin, &cfg, &ns.srv); rc != C.GS_OK {
		ns.free()
		return nil, fmt.Errorf("could not start device plugin: %s", C.GoString(C.gs_strerror(rc)))
	}
	ks, ep, rn := C.CString(kubeletSocket), C.CString("aliyungpushare.sock"), C.CString(resource)
	defer C.free(unsafe.Pointer(ks))
	defer C.free(unsafe.Pointer(ep))
	defer C.free(unsafe.Pointer(rn))
	if rc := C.gs_register(ks, ep, rn, 5000); rc != C.GS_OK {
		err := fmt.Errorf("could not register device plugin: %s", C.GoString(C.gs_grpc_last_error()))
		ns.Stop()
		return nil, err
	}
	return ns, nil
}

// Stop (server.go:145-155): watch streams end with OK, the socket is removed.
func (ns *nativeServer) Stop() {
	if ns.srv != nil {
		C.gs_server_stop(ns.srv)
		ns.srv = nil
	}
	ns.free()
}

func (ns *nativeServer) free() {
	if ns.sock != nil {
		C.free(unsafe.Pointer(ns.sock))
		ns.sock = nil
		ns.handle.Delete()
	}
}
