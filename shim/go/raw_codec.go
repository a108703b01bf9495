// raw_codec.go -- lets grpc-go hand the plugin the bytes kubelet sent, and send the bytes the
// C ABI produced, without gogo (un)marshalling in between. UNVERIFIED (no Go toolchain here).
//
// Registered on the server with grpc.ForceServerCodec(rawCodec{}); the DevicePlugin service is
// then served through a grpc.ServiceDesc whose handlers take and return *rawMessage.
package nvidia

import "fmt"

type rawMessage struct{ b []byte }

type rawCodec struct{}

func (rawCodec) Name() string { return "proto" } // what kubelet's client announces

func (rawCodec) Marshal(v interface{}) ([]byte, error) {
	m, ok := v.(*rawMessage)
	if !ok {
		return nil, fmt.Errorf("rawCodec: unexpected %T", v)
	}
	return m.b, nil
}

func (rawCodec) Unmarshal(data []byte, v interface{}) error {
	m, ok := v.(*rawMessage)
	if !ok {
		return fmt.Errorf("rawCodec: unexpected %T", v)
	}
	m.b = append(m.b[:0], data...)
	return nil
}
