"""Read an .ncu-rep (ncu --set full) and write a per-kernel summary CSV + profiles/traffic.json (DRAM bytes per ABI call = the
kernels that call launches, summed; bench.py reports it as roofline.traffic).
    python tools/summarize_ncu.py gpurun_out/prof_step.ncu-rep profiles/rNN_ncu_full_summary.csv profiles/traffic.json [steps]"""
import csv
import io
import json
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
# kernel-name fragment -> C-ABI entry point that launches it (first match wins)
ABI = [('voxel_fuse_gather', 'vfd_voxel_fuse_fwd'), ('voxel_fuse_scatter', 'vfd_voxel_fuse_bwd'),
       ('dense_dw', 'vfd_voxel_fuse_bwd'), ('voxel_fuse_reduce', 'vfd_voxel_fuse_bwd'),
       # launched by both directions: occurrences alternate forward, backward within a step (see ALTERNATE)
       ('dense_proj', 'vfd_voxel_fuse_fwd'), ('wcat_prep', 'vfd_voxel_fuse_fwd'), ('ray_sample_fwd', 'vfd_ray_sample_fwd'),
       ('ray_sample_bwd', 'vfd_ray_sample_bwd'), ('warp_stats', 'vfd_warp_stats'),
       ('warp_loss_fwd', 'vfd_warp_loss_fwd'), ('warp_loss_set_kernel', 'vfd_warp_loss_train_fwd'), ('warp_loss_bwd_kernel<1>', 'vfd_warp_loss_train_fwd'),
       ('warp_loss_bwd_kernel<true>', 'vfd_warp_loss_train_fwd'), ('warp_loss_bwd', 'vfd_warp_loss_bwd'),
       ('warp_grad_combine', 'vfd_warp_loss_train_bwd'), ('warp_dT_reduce', 'vfd_warp_loss_train_bwd'),
       ('to_depth_fwd', 'vfd_to_depth_fwd'), ('to_depth_bwd', 'vfd_to_depth_bwd'),
       ('to_depth_same_fwd', 'vfd_to_depth_fwd'), ('to_depth_same_bwd', 'vfd_to_depth_bwd'),
       ('depth_syn_fwd', 'vfd_depth_syn_fwd'), ('depth_syn_bwd', 'vfd_depth_syn_bwd'), ('voxel_count', 'vfd_voxel_classify'),
       ('voxel_scan', 'vfd_voxel_classify'), ('voxel_slot', 'vfd_voxel_classify'),
       ('smooth_mean', 'vfd_smooth_loss_train_fwd'), ('smooth_loss_kernel', 'vfd_smooth_loss_fwd'),
       ('smooth_bwd_kernel<1>', 'vfd_smooth_loss_train_fwd'), ('smooth_bwd_kernel<true>', 'vfd_smooth_loss_train_fwd'),
       ('smooth_bwd_finish', 'vfd_smooth_loss_train_bwd'), ('smooth_bwd', 'vfd_smooth_loss_bwd'), ('voxel_gather_fwd', 'vfd_voxel_gather_depth_fwd'),
       ('voxel_gather_bwd', 'vfd_voxel_gather_depth_bwd')]
ALTERNATE = {'dense_proj': ('vfd_voxel_fuse_fwd', 'vfd_voxel_fuse_bwd'), 'wcat_prep': ('vfd_voxel_fuse_fwd', 'vfd_voxel_fuse_bwd')}
UNIT = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0}


def main(rep, out_csv, out_json, steps=1):
    steps = int(steps)
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    name_i = col['Kernel Name']
    traffic = {}
    seen = {}
    with open(out_csv, 'w', newline='') as f:
        w = csv.writer(f)
        w.writerow(['Kernel Name'] + KEYS)
        w.writerow([''] + [units[col[k]] for k in KEYS])
        for r in rows[2:]:
            w.writerow([r[name_i]] + [r[col[k]] for k in KEYS])
            for frag, abi in ABI:
                if frag in r[name_i]:
                    b = sum(float(r[col[k]].replace(',', '')) * UNIT[units[col[k]]]
                            for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))
                    if frag in ALTERNATE:
                        abi = ALTERNATE[frag][seen.get(frag, 0) % 2]
                        seen[frag] = seen.get(frag, 0) + 1
                    traffic.setdefault(abi, []).append(b)
                    break
    # the capture covers `steps` steps (default 1): DRAM bytes of all kernels one ABI call launches, per call
    with open(out_json, 'w') as f:
        json.dump({k: sum(v) / steps for k, v in traffic.items()}, f, indent=1, sort_keys=True)
    print({k: len(v) for k, v in traffic.items()})
    print(open(out_csv).read())


if __name__ == '__main__':
    main(*sys.argv[1:5])
