# Builds libpdb200.so (the C-ABI product library, sm_100a only) in-tree, and the test oracle.
#   make            -> paddlepaddle_b200/libpdb200.so
#   make oracle     -> oracle/liboracle.so (+ oracle/_ref when /root/reference is present)
NVCC      ?= nvcc
NVFLAGS   := -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo --expt-relaxed-constexpr \
             -Xcompiler -fPIC -Xcompiler -fvisibility=hidden
CSRC      := paddlepaddle_b200/csrc
OBJDIR    := build
SRCS      := runtime elementwise elementwise_more elementwise_pred unary elementwise_grad reduce_grad cross_entropy rms_norm peer softmax_grad layer_norm_grad reduce_sum reduce_minmax reduce_prod argmax softmax softmax_mask layer_norm fused_ln no_alloc
OBJS      := $(addprefix $(OBJDIR)/,$(addsuffix .o,$(SRCS)))
LIB       := paddlepaddle_b200/libpdb200.so
HDRS      := $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.h) include/pdb200.h

all: $(LIB)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) -shared -o $@ $(OBJS) -lcudart_static -lpthread -ldl -lrt

oracle: oracle/liboracle.so
	./oracle/build_ref.sh

oracle/liboracle.so: oracle/phi_cpu_oracle.c oracle/phi_cpu_oracle_family.c
	gcc -O2 -fPIC -shared -std=gnu11 -fno-fast-math -ffp-contract=off -o $@ $^ -lm

clean:
	rm -rf $(OBJDIR) $(LIB) oracle/liboracle.so

.PHONY: all oracle clean

