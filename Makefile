# yalm_b200 build: one shared library (kernels + C ABI + C++ API), plus the reference-shaped CLIs.
#   make            -> yalm_b200/libyalm_b200.so
#   make cli        -> build/main (the reference's `main <ckpt> -d cuda ...` command line) and build/test (kernel
#                      benchmarks, the reference's `test -k <name>` command line)
#   make BUILD=build/x LIB=build/lib_x.so EXTRA=-DYALM_MEGA_LL=0   -> an A/B variant (YALM_B200_LIB=... selects it at run time)
NVCC  ?= nvcc
HOSTCXX := /usr/bin/g++
ARCH  := -gencode arch=compute_100a,code=sm_100a
CSRC  := yalm_b200/csrc
BUILD := build
EXTRA ?=
NVFLAGS := -std=c++17 -O3 -lineinfo $(EXTRA) $(ARCH) -ccbin $(HOSTCXX) -Xcompiler -fPIC,-Wall,-fopenmp -Iinclude
LIB   := yalm_b200/libyalm_b200.so

LIB_SRCS := $(CSRC)/ops_gemv.cu $(CSRC)/ops_attn.cu $(CSRC)/decode_mega.cu $(CSRC)/decode_mega_f32.cu $(CSRC)/decode_mega_f16.cu $(CSRC)/decode_mega_f8.cu $(CSRC)/decode_mega_f32_pipe.cu $(CSRC)/decode_mega_f16_pipe.cu $(CSRC)/decode_mega_f8_pipe.cu $(CSRC)/decode_mega_f16_mma.cu $(CSRC)/model.cpp $(CSRC)/codec.cpp $(CSRC)/capi.cpp \
            $(CSRC)/mem.cpp $(CSRC)/tp.cpp $(CSRC)/loader.cpp $(CSRC)/prefill_gemm.cu $(CSRC)/prefill.cu $(CSRC)/ops_sample.cu $(CSRC)/sampler.cpp $(CSRC)/tokenizer.cpp
LIB_OBJS := $(patsubst $(CSRC)/%,$(BUILD)/%.o,$(LIB_SRCS))
HDRS := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh include/*.h)

all: lib
cli: $(BUILD)/test $(BUILD)/main $(BUILD)/test_kernels
lib: $(LIB)

$(BUILD)/%.cu.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(BUILD)/%.cpp.o: $(CSRC)/%.cpp $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

$(LIB): $(LIB_OBJS)
	$(NVCC) $(ARCH) -ccbin $(HOSTCXX) -shared -o $@ $^ -lcudart -ldl -Xcompiler -fopenmp

$(BUILD)/test: $(CSRC)/test.cpp $(LIB) $(HDRS)
	$(NVCC) $(NVFLAGS) -x cu $(CSRC)/test.cpp -o $@ -Lyalm_b200 -lyalm_b200 \
	    -Xlinker -rpath,'$$ORIGIN/../yalm_b200'

$(BUILD)/main: $(CSRC)/main.cpp $(LIB) $(HDRS)
	$(NVCC) $(NVFLAGS) -x cu $(CSRC)/main.cpp -o $@ -Lyalm_b200 -lyalm_b200 \
	    -Xlinker -rpath,'$$ORIGIN/../yalm_b200'

# test infrastructure: a reference-style C++ caller compiled against csrc/yalm_api.h (tests/cpp), linked with the oracle
$(BUILD)/test_kernels: tests/cpp/test_kernels.cpp $(LIB) $(HDRS) oracle/liboracle.so
	$(NVCC) -std=c++17 -O1 $(ARCH) -ccbin $(HOSTCXX) -x cu tests/cpp/test_kernels.cpp -I$(CSRC) -Iinclude -o $@ \
	    -Lyalm_b200 -lyalm_b200 -Loracle -loracle -Xlinker -rpath,'$$ORIGIN/../yalm_b200' -Xlinker -rpath,'$$ORIGIN/../oracle'

oracle/liboracle.so: oracle/yalm_oracle.c oracle/yalm_oracle.h
	$(MAKE) -s -C oracle liboracle.so

clean:
	rm -rf $(BUILD) $(LIB)

.PHONY: all lib cli clean
