# Builds the B200-native d2q9-bgk: liblbm_b200.so (CUDA, sm_100a) + the C host driver.
# Same targets as the reference Makefile (all / check / clean), same exe name.
EXE = d2q9-bgk
PKG = lattice-boltzmann_b200
LIB = $(PKG)/liblbm_b200.so

NVCC ?= nvcc
CC = gcc
NVFLAGS = -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC
CFLAGS = -std=c99 -Wall -O2 -fopenmp

FINAL_STATE_FILE = ./final_state.dat
AV_VELS_FILE = ./av_vels.dat
REF_FINAL_STATE_FILE = check/128x128.final_state.dat
REF_AV_VELS_FILE = check/128x128.av_vels.dat
# the reference's own checker when its tree is present, else our restatement of it
CHECKER = $(if $(wildcard /root/reference/check/check.py),/root/reference/check/check.py,tools/check.py)

all: $(EXE)

$(LIB): $(PKG)/csrc/lbm_api.cu $(PKG)/csrc/lbm_plan.cpp $(PKG)/csrc/lbm_kernels.cuh $(PKG)/csrc/lbm_device.cuh $(PKG)/csrc/lbm_plan.hpp include/lbm_b200.h
	$(NVCC) $(NVFLAGS) -shared $(PKG)/csrc/lbm_api.cu $(PKG)/csrc/lbm_plan.cpp -o $@

# the self-checking build (stand-in for compute-sanitizer, see lbm_device.cuh): LBM_LIBRARY=$(LIB_CHECKED) pytest -m gpu
LIB_CHECKED = $(PKG)/liblbm_b200_checked.so
checked: $(LIB_CHECKED)
$(LIB_CHECKED): $(PKG)/csrc/lbm_api.cu $(PKG)/csrc/lbm_plan.cpp $(PKG)/csrc/lbm_kernels.cuh $(PKG)/csrc/lbm_device.cuh $(PKG)/csrc/lbm_plan.hpp include/lbm_b200.h
	$(NVCC) $(NVFLAGS) -DLBM_CHECKED -shared $(PKG)/csrc/lbm_api.cu $(PKG)/csrc/lbm_plan.cpp -o $@

$(EXE): src/d2q9-bgk.c src/fmt_e12.h include/lbm_b200.h $(LIB)
	$(CC) $(CFLAGS) src/d2q9-bgk.c -L$(PKG) -llbm_b200 -Wl,-rpath,'$$ORIGIN/$(PKG)' -lm -o $@

oracle:
	$(MAKE) -C oracle

# the known answers are regenerated from the committed fixtures (tests/golden/shipped_*.npz) on first use
$(REF_FINAL_STATE_FILE) $(REF_AV_VELS_FILE):
	python tools/make_check_files.py

check: $(REF_FINAL_STATE_FILE) $(REF_AV_VELS_FILE)
	python $(CHECKER) --ref-av-vels-file=$(REF_AV_VELS_FILE) --ref-final-state-file=$(REF_FINAL_STATE_FILE) --av-vels-file=$(AV_VELS_FILE) --final-state-file=$(FINAL_STATE_FILE)

# the parity suite against the self-checking build (GPU box); compute-sanitizer itself is closed on this pool
sanitize: all checked
	LBM_LIBRARY=$(CURDIR)/$(LIB_CHECKED) python -m pytest tests -m gpu -q -k "oracle or slabs or engines or checked"

.PHONY: all check clean oracle sanitize checked

clean:
	rm -f $(EXE) $(LIB) $(LIB_CHECKED)
