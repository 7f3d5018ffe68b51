"""CPU tests of the profiling helpers under tools/ (no GPU, no native library)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CSV = '''==PROF== Connected to process 1
"ID","Process ID","Process Name","Host Name","Kernel Name","Context","Stream","Block Size","Grid Size","Device","CC","Section Name","Metric Name","Metric Unit","Metric Value"
"0","1","python","h","void pmd::<unnamed>::pad_stem_weight_kernel(const __nv_bfloat16 *, __nv_bfloat16 *, int, int)","1","14","(256, 1, 1)","(16, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","2900"
"1","1","python","h","void pmd::<unnamed>::augment_im2col_kernel(const unsigned char *, const int *, int, int, int, int, __nv_bfloat16 *)","1","14","(128, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","31000"
"2","1","python","h","void pmd::<unnamed>::conv_gemm_kernel<64, 2, 0, 0>(pmd::ConvGemmParams)","1","14","(192, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","48,000"
"3","1","python","h","void pmd::<unnamed>::conv_gemm_kernel<64, 2, 0, 0>(pmd::ConvGemmParams)","1","14","(192, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","52000"
"4","1","python","h","void pmd::<unnamed>::pad_stem_weight_kernel(const __nv_bfloat16 *, __nv_bfloat16 *, int, int)","1","14","(256, 1, 1)","(16, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","3000"
"5","1","python","h","void pmd::<unnamed>::augment_im2col_kernel(const unsigned char *, const int *, int, int, int, int, __nv_bfloat16 *)","1","14","(128, 1, 1)","(4096, 1, 1)","0","10.0","Command line profiler metrics","gpu__time_duration.sum","ns","30000"
'''


def test_launch_table_picks_the_last_complete_step(tmp_path):
    src = tmp_path / 'l.csv'
    src.write_text(CSV)
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'launch_table.py'), str(src), 'unit test'],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    out = r.stdout
    assert out.startswith('# unit test')
    # the step is launches 1..4 (augment .. pad_stem); the trailing augment without a closing pad is not a step
    assert '| conv_gemm_kernel<64, 2, 0, 0> | 2 | 100.0 |' in out
    assert '| **total (serialised)** | 4 | 134.0 | |' in out
    assert '| 0 | augment_im2col_kernel | (4096, 1, 1) | 31.0 |' in out
