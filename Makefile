# Build of the B200 lab suite: one native library (shared for Python, static for
# the CLI binaries) plus the contract binaries at the paths the harness expects
# (<labN>/src/<exe>, ref: run_test.py:58-60, README.md:36-42).
NVCC      ?= nvcc
CXX       ?= g++
CC        ?= gcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -lineinfo -std=c++17 -Iinclude -Icsrc -Xcompiler -fPIC \
             -Werror cross-execution-space-call $(EXTRA_NVFLAGS)
# oracles must not be contracted by the host compiler: FMAs are spelled out
CXXFLAGS  := -O2 -std=c++17 -fPIC -Iinclude -ffp-contract=off
CFLAGS    := -O2 -std=c11
BUILD     := build
LIBDIR    := cuda_mpi_openmp_b200/lib

CU_SRCS   := $(wildcard csrc/*.cu)
CU_OBJS   := $(patsubst csrc/%.cu,$(BUILD)/%.o,$(CU_SRCS))
CPP_OBJS  := $(BUILD)/oracle.o
# the stand-alone CPU binaries get the multi-threaded build of the oracles (see csrc/oracle.cpp)
MT_OBJS  := $(BUILD)/oracle_mt.o
HDRS      := $(wildcard csrc/*.cuh) include/b200labs.h $(wildcard apps/*.h)

SO        := $(LIBDIR)/libb200labs.so
AR_LIB    := $(BUILD)/libb200labs.a

CUDA_BINS := lab1/src/cuda_exe lab1/src/cuda_exe_to_plot \
             lab2/src/cuda_exe lab2/src/to_plot_cuda_exe \
             lab3/src/cuda_exe lab3/src/to_plot_cuda_exe \
             lab5/src/cuda_exe_int lab5/src/cuda_exe_float lab5/src/cuda_exe_uchar \
             lab5/src/to_plot_cuda_exe gpu_info/src/gpu_info
CPU_BINS  := lab1/src/cpu_exe lab1/src/cpu_exe_to_plot lab2/src/cpu_exe lab3/src/cpu_exe \
             lab3/src/test_read_input lab5/src/cpu_exe_int lab5/src/cpu_exe_float \
             lab5/src/cpu_exe_uchar lab5/src/to_plot_cpu_exe hw1/src/hw1 hw2/src/hw2

# the binaries' directories hold nothing that is tracked, so a fresh checkout lacks them
$(shell mkdir -p lab1/src lab2/src lab3/src lab5/src hw1/src hw2/src gpu_info/src)

.PHONY: all lib bins cpu clean sass yardsticks
all: lib bins
lib: $(SO) $(AR_LIB)
bins: $(CUDA_BINS) $(CPU_BINS)
cpu: $(CPU_BINS)

$(BUILD)/%.o: csrc/%.cu $(HDRS) | $(BUILD)
	$(NVCC) $(NVFLAGS) -c $< -o $@
$(BUILD)/oracle.o: csrc/oracle.cpp include/b200labs.h | $(BUILD)
	$(CXX) $(CXXFLAGS) -c $< -o $@
$(BUILD)/oracle_mt.o: csrc/oracle.cpp include/b200labs.h | $(BUILD)
	$(CXX) $(CXXFLAGS) -DB200_ORACLE_THREADS -pthread -c $< -o $@
$(BUILD):
	mkdir -p $(BUILD)

$(SO): $(CU_OBJS) $(CPP_OBJS)
	mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -o $@ $^
$(AR_LIB): $(CU_OBJS) $(CPP_OBJS)
	rm -f $@ && ar rcs $@ $^

# ---- CUDA contract binaries (statically linked against the library) ----------
APP_NV := $(NVCC) $(NVFLAGS) -Iapps
lab1/src/cuda_exe: apps/lab1_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) $< $(AR_LIB) -o $@
lab1/src/cuda_exe_to_plot: apps/lab1_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) -DTO_PLOT $< $(AR_LIB) -o $@
lab2/src/cuda_exe: apps/lab2_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) $< $(AR_LIB) -o $@
lab2/src/to_plot_cuda_exe: apps/lab2_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) -DTO_PLOT $< $(AR_LIB) -o $@
lab3/src/cuda_exe: apps/lab3_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) $< $(AR_LIB) -o $@
lab3/src/to_plot_cuda_exe: apps/lab3_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) -DTO_PLOT $< $(AR_LIB) -o $@
lab5/src/cuda_exe_int: apps/lab5_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) -DKEY_TYPE=0 $< $(AR_LIB) -o $@
lab5/src/cuda_exe_float: apps/lab5_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) -DKEY_TYPE=1 $< $(AR_LIB) -o $@
lab5/src/cuda_exe_uchar: apps/lab5_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) -DKEY_TYPE=2 $< $(AR_LIB) -o $@
lab5/src/to_plot_cuda_exe: apps/lab5_cuda.cu $(AR_LIB) $(HDRS)
	$(APP_NV) -DTO_PLOT $< $(AR_LIB) -o $@
gpu_info/src/gpu_info: apps/gpu_info.cu $(AR_LIB) $(HDRS)
	$(APP_NV) $< $(AR_LIB) -o $@

# ---- CPU binaries (no CUDA needed: `make cpu` works on any box) ---------------
APP_CXX := $(CXX) $(CXXFLAGS) -pthread -Iapps
lab1/src/cpu_exe: apps/lab1_cpu.cpp $(MT_OBJS)
	$(APP_CXX) -DNO_TIME_LINE $^ -o $@
lab1/src/cpu_exe_to_plot: apps/lab1_cpu.cpp $(MT_OBJS)
	$(APP_CXX) $^ -o $@
lab2/src/cpu_exe: apps/lab2_cpu.cpp $(MT_OBJS)
	$(APP_CXX) $^ -o $@
lab3/src/cpu_exe: apps/lab3_cpu.cpp $(MT_OBJS)
	$(APP_CXX) $^ -o $@
lab3/src/test_read_input: apps/lab3_read_input.c
	$(CC) $(CFLAGS) $< -o $@
lab5/src/cpu_exe_int: apps/lab5_cpu.cpp $(MT_OBJS)
	$(APP_CXX) -DKEY_TYPE=0 $^ -o $@
lab5/src/cpu_exe_float: apps/lab5_cpu.cpp $(MT_OBJS)
	$(APP_CXX) -DKEY_TYPE=1 $^ -o $@
lab5/src/cpu_exe_uchar: apps/lab5_cpu.cpp $(MT_OBJS)
	$(APP_CXX) -DKEY_TYPE=2 $^ -o $@
lab5/src/to_plot_cpu_exe: apps/lab5_cpu.cpp $(MT_OBJS)
	$(APP_CXX) -DTO_PLOT $^ -o $@
hw1/src/hw1: apps/hw1.c
	$(CC) $(CFLAGS) $< -lm -o $@
hw2/src/hw2: apps/hw2.c
	$(CC) $(CFLAGS) $< -o $@

# Yardsticks and reference drivers (measurement only; nothing here is linked into the library)
yardsticks: | $(BUILD)
	$(NVCC) -O3 -std=c++17 -arch=sm_100a -o $(BUILD)/sort_yardstick tools/sort_yardstick.cu
	bash baseline/build_lab1_driver.sh

# SASS listings for profiles/ (evidence of UTMALDG/UTMASTG, vector widths, ...)
sass: $(SO)
	mkdir -p profiles/sass
	for o in $(CU_OBJS); do cuobjdump -sass $$o | sed -e 's#/\* 0x[0-9a-f]* \*/##' -e 's/[ \t]*$$//' | grep -v '^$$' \
	    > profiles/sass/$$(basename $$o .o).sass; done

clean:
	rm -rf $(BUILD) $(SO) $(CUDA_BINS) $(CPU_BINS)
