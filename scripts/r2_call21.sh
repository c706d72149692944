#!/bin/bash
# L1 eviction hints on node / primitive loads, shade: prefetch of the next entry's primitive record, st.global.cg state stores
mkdir -p gpurun_out
bash scripts/ab2.sh "base;;raytracer-cuda_b200/libcrt_b200.so" "ldcg2;;raytracer-cuda_b200/build/var/libldcg2.so" "nodeEL;;raytracer-cuda_b200/build/var/libnodeEL.so" "nodeELprimEF;;raytracer-cuda_b200/build/var/libnodeELprimEF.so" "primNA;;raytracer-cuda_b200/build/var/libprimNA.so" "shpf(ldcg2+prim prefetch);;raytracer-cuda_b200/build/var/libshpf.so" "stcg(ldcg2+st.cg);;raytracer-cuda_b200/build/var/libstcg.so" > gpurun_out/d9_ab.log 2>&1
cat gpurun_out/d9_ab.log
