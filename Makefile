# Build of libhmj.so (CUDA, sm_100a only), the CPU oracle, the reference shim and the C++
# facade test binaries.  `make` = everything that can be built on this machine.
NVCC      ?= nvcc
CXX       ?= g++
CC        ?= gcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -Wall
CSRC      := hashmergejoin_b200/csrc
BUILD     := build
LIB       := hashmergejoin_b200/libhmj.so
CU_SRCS   := $(wildcard $(CSRC)/*.cu)
CU_OBJS   := $(patsubst $(CSRC)/%.cu,$(BUILD)/%.o,$(CU_SRCS))
CU_HDRS   := $(wildcard $(CSRC)/*.cuh) include/hmj.h
REFERENCE ?= /root/reference

.PHONY: all lib oracle ref tests clean
all: lib oracle ref tests

lib: $(LIB)

$(BUILD)/%.o: $(CSRC)/%.cu $(CU_HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(CU_OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart_static -lpthread -ldl -lrt

# ---- CPU oracle (test infrastructure only) -------------------------------------------
oracle: oracle/libhmj_oracle.so
oracle/libhmj_oracle.so: oracle/hmj_oracle.c oracle/hmj_oracle.h
	$(CC) -O2 -std=c11 -Wall -fPIC -shared -o $@ oracle/hmj_oracle.c -lm

# ---- the reference itself, compiled from where it lies (only when it is mounted) ------
ref:
	@if [ -d $(REFERENCE) ]; then $(MAKE) oracle/_ref/libhmj_ref.so; \
	 else echo "reference not mounted: keeping prebuilt oracle/_ref (if any)"; fi
oracle/_ref/libhmj_ref.so: oracle/ref_shim.cc
	@mkdir -p oracle/_ref
	$(CXX) -O2 -std=c++11 -pthread -fPIC -shared -include cstdint -I$(REFERENCE) \
	    -o $@ oracle/ref_shim.cc $(REFERENCE)/thread_barrier.cc

# ---- C++ facade tests (mirror the reference's gtest files against the facade headers) --
FACADE    := hashmergejoin_b200/facade
TEST_BINS := $(BUILD)/facade_test $(BUILD)/facade_bench $(BUILD)/facade_bench_borrow
tests: $(TEST_BINS)
$(BUILD)/facade_test: tests/cpp/facade_test.cc $(wildcard $(FACADE)/*.h) include/hmj.h $(LIB)
	@mkdir -p $(BUILD)
	$(CXX) -O2 -std=c++11 -pthread -Iinclude -I$(FACADE) -o $@ tests/cpp/facade_test.cc \
	    -Lhashmergejoin_b200 -lhmj -Wl,-rpath,'$$ORIGIN/../hashmergejoin_b200'

$(BUILD)/facade_bench: tests/cpp/facade_bench.cc $(wildcard $(FACADE)/*.h) include/hmj.h $(LIB)
	@mkdir -p $(BUILD)
	$(CXX) -O2 -std=c++11 -pthread -Iinclude -I$(FACADE) -o $@ tests/cpp/facade_bench.cc \
	    -Lhashmergejoin_b200 -lhmj -Wl,-rpath,'$$ORIGIN/../hashmergejoin_b200'

$(BUILD)/facade_bench_borrow: tests/cpp/facade_bench.cc $(wildcard $(FACADE)/*.h) include/hmj.h $(LIB)
	@mkdir -p $(BUILD)
	$(CXX) -O2 -std=c++11 -pthread -DHMJ_FACADE_BORROW_INPUTS -Iinclude -I$(FACADE) -o $@ tests/cpp/facade_bench.cc \
	    -Lhashmergejoin_b200 -lhmj -Wl,-rpath,'$$ORIGIN/../hashmergejoin_b200'

clean:
	rm -rf $(BUILD) $(LIB) oracle/libhmj_oracle.so oracle/_ref

# ---- instrumented build (per-tile timeline of the scatter pass; scripts/trace_pass.py) ----
TRACE_OBJS := $(patsubst $(CSRC)/%.cu,$(BUILD)/trace/%.o,$(CU_SRCS))
$(BUILD)/trace/%.o: $(CSRC)/%.cu $(CU_HDRS)
	@mkdir -p $(BUILD)/trace
	$(NVCC) $(NVFLAGS) -DHMJ_TRACE -c $< -o $@
trace: $(BUILD)/trace/libhmj.so
$(BUILD)/trace/libhmj.so: $(TRACE_OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart_static -lpthread -ldl -lrt
