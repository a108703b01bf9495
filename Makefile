# Builds libgpushare_core.so (product), the oracle/test helpers and the fake NVML.
# Everything lands in-tree so the built files travel to the GPU box with gpurun.
NVCC      ?= nvcc
CXX       ?= g++
CC        ?= gcc
PKG       := aliyun-gpushare-device-plugin_b200
CSRC      := $(PKG)/csrc
ARCH      := -gencode arch=compute_100a,code=sm_100a
CUDA_INC  ?= /usr/local/cuda/include
NVCCFLAGS := $(ARCH) -lineinfo -O3 -std=c++17 -Iinclude -I$(CSRC) -Xcompiler -fPIC,-fvisibility=hidden,-Wall
CXXFLAGS  := -O2 -g -std=c++17 -fPIC -fvisibility=hidden -Wall -Wextra -Iinclude -I$(CSRC) -I$(CUDA_INC)
CFLAGS    := -O2 -g -std=gnu11 -fPIC -Wall -Wextra

CORE_SO   := $(PKG)/libgpushare_core.so
CPP_SRCS  := $(wildcard $(CSRC)/*.cpp)
CU_SRCS   := $(wildcard $(CSRC)/*.cu)
OBJS      := $(CPP_SRCS:.cpp=.o) $(CU_SRCS:.cu=.o)

ORACLE_SO := oracle/libgs_oracle.so
FAKE_NVML := tests/fake_nvml/libfake_nvml.so

DAEMON    := $(PKG)/bin/gpushare-device-plugin-v2
INSPECT   := $(PKG)/bin/kubectl-inspect-gpushare-v2

all: $(CORE_SO) $(ORACLE_SO) $(FAKE_NVML) $(DAEMON) $(INSPECT)

$(CSRC)/%.o: $(CSRC)/%.cpp include/gpushare.h $(wildcard $(CSRC)/*.h)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(CSRC)/%.o: $(CSRC)/%.cu include/gpushare.h $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@

$(CORE_SO): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -ldl -lpthread

# The node daemon as a native binary (flags, kube I/O over OpenSSL, supervisor loop) over the core library.
$(DAEMON): $(wildcard $(PKG)/daemon_cpp/*.cpp $(PKG)/daemon_cpp/*.h) include/gpushare.h $(CORE_SO)
	mkdir -p $(PKG)/bin
	$(CXX) -O2 -g -std=c++17 -Wall -Wextra -Iinclude -I$(PKG)/daemon_cpp $(PKG)/daemon_cpp/main.cpp -o $@ \
		-L$(PKG) -lgpushare_core -Wl,-rpath,'$$ORIGIN/..' -lssl -lcrypto -lpthread

$(INSPECT): $(wildcard $(PKG)/daemon_cpp/*.cpp $(PKG)/daemon_cpp/*.h) include/gpushare.h $(CORE_SO)
	mkdir -p $(PKG)/bin
	$(CXX) -O2 -g -std=c++17 -Wall -Wextra -Iinclude -I$(PKG)/daemon_cpp $(PKG)/daemon_cpp/inspect_main.cpp -o $@ \
		-L$(PKG) -lgpushare_core -Wl,-rpath,'$$ORIGIN/..' -lssl -lcrypto -lpthread

# Oracle: CPU restatement used ONLY by tests/, smoke() and bench.py's cpu_baseline leg.
ORACLE_OBJS := $(patsubst %.c,%.o,$(wildcard oracle/*.c)) $(patsubst %.cpp,%.o,$(wildcard oracle/*.cpp))
oracle/%.o: oracle/%.c
	$(CC) $(CFLAGS) -c $< -o $@
oracle/%.o: oracle/%.cpp $(wildcard oracle/*.h)
	$(CXX) -O2 -g -std=c++17 -fPIC -Wall -Wextra -c $< -o $@
$(ORACLE_SO): $(ORACLE_OBJS)
	$(CXX) -shared -o $@ $(ORACLE_OBJS) -lpthread

$(FAKE_NVML): tests/fake_nvml/fake_nvml.c
	$(CC) $(CFLAGS) -shared -o $@ $< -lpthread

clean:
	rm -f $(OBJS) $(CORE_SO) $(ORACLE_OBJS) $(ORACLE_SO) $(FAKE_NVML) $(DAEMON) $(INSPECT)

test: all
	python -m pytest tests -q -m "not gpu"

test-gpu: all
	python -m pytest tests -q -m gpu

.PHONY: all clean test test-gpu
