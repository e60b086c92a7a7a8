# Builds the product library (CUDA, sm_100a only) and the CPU oracle (test infrastructure).
NVCC ?= nvcc
CSRC := constraint_solver_b200/csrc
LIB  := constraint_solver_b200/lib/libxpbd_b200.so
NVCCFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false \
             -Xcompiler -fPIC,-ffp-contract=off,-Wall -shared
SRCS := $(wildcard $(CSRC)/*.cu) $(wildcard $(CSRC)/*.cpp)
HDRS := $(wildcard $(CSRC)/*.cuh $(CSRC)/*.hpp) include/xpbd_b200.h

all: lib oracle
lib: $(LIB)
$(LIB): $(SRCS) $(HDRS)
	mkdir -p $(dir $(LIB))
	$(NVCC) $(NVCCFLAGS) $(EXTRA) -o $@ $(SRCS)
# debug build that counts how many edge pairs survive the SAT stage's witness cull (scripts/cull_stats.py)
stats:
	mkdir -p constraint_solver_b200/lib/debug
	$(NVCC) $(NVCCFLAGS) -DXPBD_CULL_STATS -o constraint_solver_b200/lib/debug/libxpbd_b200_stats.so $(SRCS)
# tuning variant: make variant NAME=sat5 EXTRA="-DXPBD_SAT_THREAD_MIN_BLOCKS=5" -> lib/variants/libxpbd_b200_sat5.so (XPBD_B200_LIB selects it)
variant:
	mkdir -p constraint_solver_b200/lib/variants
	$(NVCC) $(NVCCFLAGS) $(EXTRA) -o constraint_solver_b200/lib/variants/libxpbd_b200_$(NAME).so $(SRCS)
oracle:
	$(MAKE) -C oracle
clean:
	rm -rf constraint_solver_b200/lib oracle/_build
.PHONY: all lib oracle clean stats variant
