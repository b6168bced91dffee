timeout 300 ncu --set full --clock-control none --import-source on -k regex:hash_bwd_table -s 27 -c 1 -o gpurun_out/hash_bwd_table_v10 -f python tools/smooth_probe.py tum_desk > gpurun_out/ncu_hbt.log 2>&1
ncu -i gpurun_out/hash_bwd_table_v10.ncu-rep --page details > gpurun_out/ncu_full_v10_hash_bwd_table.txt 2>&1
grep -E "Duration|DRAM Throughput|Memory Throughput|L2 Cache Throughput|Registers Per|Achieved Occupancy|Issue Slots Busy|L1/TEX Hit|L2 Hit|Theoretical Occupancy|Executed Ipc" gpurun_out/ncu_full_v10_hash_bwd_table.txt | head -20
