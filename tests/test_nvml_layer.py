"""NVML layer against the fake libnvidia-ml (BASELINE configs 1 and 5)."""
import ctypes as C
import threading
import time

import pytest

from gpushare_b200 import api
from oracle import ref_model as R
from tests.conftest import FAKE_NVML


@pytest.fixture
def fake_env(monkeypatch):
    def setup(gpus=8, mem_mib=16384, no_events=""):
        monkeypatch.setenv("FAKE_NVML_GPUS", str(gpus))
        monkeypatch.setenv("FAKE_NVML_MEM_MIB", str(mem_mib))
        monkeypatch.setenv("FAKE_NVML_NO_EVENTS", no_events)
    return setup


def test_enumerate_config1(fake_env):
    fake_env(8, 16384)
    with api.Nvml(FAKE_NVML) as nv:
        assert nv.device_count() == 8
        infos = nv.devices()
        assert infos[3].uuid.decode() == "GPU-00000003-0000-0000-0000-000000000003"
        assert infos[3].total_bytes == 16384 << 20 and infos[3].nvml_index == 3
        assert infos[3].pci_bus_id.decode() == "00000000:13:00.0"
        t = api.DeviceTable.from_gpus(infos, "GiB")
        assert (t.gpu_count, t.units_per_gpu, t.device_count) == (8, 16, 128)
        g = R.Globals("GiB")
        devs, _ = R.getDevices(g, [(i.uuid.decode(), i.total_bytes >> 20) for i in infos])
        assert t.list_and_watch_bytes() == R.encode_list_and_watch(devs)
        t2 = api.DeviceTable.from_gpus(infos, "MiB")
        assert t2.units_per_gpu == 16384


def test_open_errors(fake_env):
    with pytest.raises(api.L.GpushareError) as e:
        api.Nvml("/nonexistent/libnvidia-ml.so.1")
    assert e.value.status == api.L.GS_E_NOLIB
    fake_env(2)
    with api.Nvml(FAKE_NVML) as nv:
        with pytest.raises(api.L.GpushareError) as e:
            api.Nvml(FAKE_NVML)                      # already open
        assert e.value.status == api.L.GS_E_STATE
        with pytest.raises(api.L.GpushareError):
            nv.device_info(5)


def test_xid_events_follow_reference_rule(fake_env):
    """Config 5 script: XID 79 on GPU 3, XID 31 (ignored) on GPU 1, an event without a
    device identity (=> all GPUs)."""
    fake_env(8, 16384, no_events="6")
    fake = C.CDLL(FAKE_NVML)
    fake.fake_nvml_push_event.argtypes = [C.c_int, C.c_ulonglong]
    with api.Nvml(FAKE_NVML) as nv:
        assert nv.health_open() == 1 << 6            # GPU 6 cannot register => unhealthy at once
        assert nv.health_wait(50) is None
        fake.fake_nvml_push_event(3, 79)
        fake.fake_nvml_push_event(1, 31)
        fake.fake_nvml_push_event(-1, 48)
        assert nv.health_wait(1000) == (3, 79, False)
        assert nv.health_wait(1000) == (1, 31, True)
        assert nv.health_wait(1000) == (-1, 48, False)
        nv.health_close()


def test_health_thread_readvertises_once_per_event(fake_env):
    fake_env(4, 4096)
    fake = C.CDLL(FAKE_NVML)
    fake.fake_nvml_push_event.argtypes = [C.c_int, C.c_ulonglong]
    with api.Nvml(FAKE_NVML) as nv:
        t = api.DeviceTable.from_gpus(nv.devices(), "GiB")
        plug = api.NvidiaDevicePlugin(t, "n")
        nv.health_open()
        msgs = []
        consumer = threading.Thread(
            target=lambda: [msgs.append(R.decode_list_and_watch(m)) for m in plug.ListAndWatch(20)])
        consumer.start()

        def pump():                                   # healthcheck() of server.go:211-229
            while not plug.stopped:
                ev = nv.health_wait(50)
                if ev and not ev[2]:
                    plug.unhealthy(ev[0])
        pumper = threading.Thread(target=pump)
        pumper.start()
        time.sleep(0.1)
        t0 = time.time()
        fake.fake_nvml_push_event(2, 79)
        while len(msgs) < 2 and time.time() - t0 < 2:
            time.sleep(0.005)
        latency = time.time() - t0
        fake.fake_nvml_push_event(0, 45)              # application error: no re-advertise
        time.sleep(0.2)
        plug.Stop()
        consumer.join(2)
        pumper.join(2)
        nv.health_close()
    assert len(msgs) == 2 and latency < 1.0
    assert [h for i, h in msgs[1] if i.startswith("GPU-00000002")] == ["Unhealthy"] * 4
    assert sum(h == "Unhealthy" for _, h in msgs[1]) == 4
