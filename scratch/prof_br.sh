# phase clocks inside the wide blind-rotation kernel (thread 0 of CTA 0).  Build first (no GPU needed):
#   cd eoc_tfhe_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=default,-O2 \
#      -DB200_BR_PROFILE -x cu -c kernels.cu -o ../../scratch/kernels_prof.o && nvcc -shared -o ../../scratch/libb200tfhe_prof.so \
#      ../../scratch/kernels_prof.o abi.o circuit.o wire.o eoc_abi.o -gencode arch=compute_100a,code=sm_100a -Xlinker -Bsymbolic -lcudart_static -lpthread -ldl -lrt
B200TFHE_LIB=$PWD/scratch/libb200tfhe_prof.so python scratch/prof_one.py ${1:-8880} 80 2
