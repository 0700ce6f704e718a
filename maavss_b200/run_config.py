"""Mirror of the reference flag set (run_config.py:4-51): same names, defaults and (quirky) types."""
import argparse


def build_parser():
    p = argparse.ArgumentParser()
    p.add_argument('-b', '--batch_size', type=int, default=4, metavar='N')
    p.add_argument('-lr', '--learning_rate', type=float, default=1e-5)
    p.add_argument('-lc', '--loss_coeff', type=float, default=0.001)
    p.add_argument('-e', '--epochs', type=int, default=10)
    p.add_argument('-s', '--steps_per_epoch', type=int, default=50)
    p.add_argument('-v', '--val_steps', type=int, default=8)
    p.add_argument('--mode_freq', type=int, default=1)
    p.add_argument('--objective_zeros', type=bool, default=False)
    p.add_argument('--data_path', type=str, default="data/raw")
    p.add_argument('--num_frames', type=int, default=8)
    p.add_argument('--num_seq', type=int, default=4)
    p.add_argument('--frame_hop', type=int, default=2)
    p.add_argument('--framerate', type=int, default=30)
    p.add_argument('--framesize', type=int, default=256)
    p.add_argument('--p_size', type=int, default=64)
    p.add_argument('--autocontrast', type=bool, default=False)
    p.add_argument('--attn_diff', type=bool, default=False)
    p.add_argument('--compress_audio', action="store_true")
    p.add_argument('--fft_len', type=int, default=256)
    p.add_argument('-a', '--hops_per_frame', type=int, default=8)
    p.add_argument('--samplerate', type=int, default=16000)
    p.add_argument('--normalize_fft', type=bool, default=True)
    p.add_argument('--normalize_output_fft', type=bool, default=False)
    p.add_argument('--use_polar', type=bool, default=False)
    p.add_argument('--noise_scalar', type=float, default=0.1)
    p.add_argument('--fc_size', type=int, default=4096)
    p.add_argument('--latent_chan', type=int, default=64)
    p.add_argument('--cb_freq', type=int, default=100)
    p.add_argument('--max_clip_len', type=int, default=None)
    p.add_argument('--split', type=float, default=0.8)
    p.add_argument('--saved_model', type=str, default=None)
    p.add_argument('--checkpoint', type=str, default=None)
    p.add_argument('--cp_dir', type=str, default="checkpoints/")
    p.add_argument('--cp_load_opt', action="store_true")
    p.add_argument('-c', action="store_true")
    p.add_argument('--no_save', action="store_true")
    p.add_argument('--cp_freq', type=int, default=0)
    return p


def model_args(argv=None):
    return build_parser().parse_args(argv)
