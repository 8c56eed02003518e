"""``python -m tacotron_prosody_multi_gpu_b200.TacotronModel.synthesize`` -- the reference's
TacotronModel/synthesize.py CLI (same flags :112-127, same sentences.txt handling :129-141, same output
tree :69-86 / :30-56) driving the B200 engine.  Checkpoint discovery through
``tf.train.get_checkpoint_state`` (:93,106) is replaced by the newest ``*.npz`` weight container in
``--checkpoint`` (random init with the reference's seed when there is none).  Launched under torchrun,
every rank synthesises its shard of the batches on its own GPU (no collective) and writes its own rows
of ``map.txt`` (rank 0 merges)."""
import argparse
import glob
import os

from ..Utils.Hyperparams import hparams
from ..Utils.Infolog import log
from ..Utils.Tacotron_synthesizer import Synthesizer


def _find_checkpoint(checkpoint):
    if checkpoint and os.path.isfile(checkpoint):
        return checkpoint
    cands = sorted(glob.glob(os.path.join(checkpoint or "", "*.npz")), key=os.path.getmtime)
    return cands[-1] if cands else None


def _rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def run_inference(checkpoint_path, output_dir, hparams, sentences):
    inference_dir = os.path.join(output_dir, 'inference')
    log_dir = os.path.join(output_dir, 'logs-inference')
    for d in (inference_dir, log_dir, os.path.join(log_dir, 'wavs'), os.path.join(log_dir, 'plots')):
        os.makedirs(d, exist_ok=True)
    log('running inference..')
    synth = Synthesizer()
    synth.load(checkpoint_path, hparams, GTA=False)
    bs = hparams.tacotron_synthesis_batch_size
    batches = [sentences[i: i + bs] for i in range(0, len(sentences), bs)]
    rank, world = _rank_world()
    map_name = 'map.txt' if world == 1 else 'map.rank%d.txt' % rank
    with open(os.path.join(inference_dir, map_name), 'w', encoding='utf-8') as file:
        for i, text in enumerate(batches):
            if i % world != rank:
                continue
            basenames = ['batch_{}_sentence_{}'.format(i, j) for j in range(len(text))]
            mel_filename = synth.synthesize(text, basenames, inference_dir, log_dir, None)
            file.write('{}|{}\n'.format(text, mel_filename))
    _merge_maps(inference_dir, rank, world)
    log('synthesized mel spectrograms of "{}" at {}'.format(batches, inference_dir))
    return inference_dir


def run_synthesis(args, checkpoint_path, output_dir, hparams):
    GTA = (args.GTA == 'True')
    synth_dir = os.path.join(output_dir, 'gta' if GTA else 'natural')
    os.makedirs(synth_dir, exist_ok=True)
    metadata_filename = os.path.join(args.input_dir, 'train.txt')
    log(hparams_debug())
    synth = Synthesizer()
    synth.load(checkpoint_path, hparams, GTA=GTA)
    with open(metadata_filename, encoding='utf-8') as f:
        metadata = [line.strip().split('|') for line in f]
    hours = sum(int(x[4]) for x in metadata) * (hparams.hop_size / hparams.sample_rate) / 3600
    log('Loaded metadata for {} examples ({:.2f} hours)'.format(len(metadata), hours))
    chunk = 64 * hparams.tacotron_num_gpus   # one persistent-kernel launch covers <= 64 utterances per tower
    metadata = [metadata[i: i + chunk] for i in range(0, len(metadata), chunk)]
    log('starting synthesis..')
    mel_dir = os.path.join(args.input_dir, 'mels')
    wav_dir = os.path.join(args.input_dir, 'audio')
    rank, world = _rank_world()
    map_name = 'map.txt' if world == 1 else 'map.rank%d.txt' % rank
    with open(os.path.join(synth_dir, map_name), 'w', encoding='utf-8') as file:
        for i, meta in enumerate(metadata):
            if i % world != rank:
                continue
            texts = [m[5] for m in meta]
            mel_filenames = [os.path.join(mel_dir, m[1]) for m in meta]
            wav_filenames = [os.path.join(wav_dir, m[0]) for m in meta]
            basenames = [os.path.basename(m).replace('.npy', '').replace('mel-', '') for m in mel_filenames]
            mel_output_filenames, speaker_ids = synth.synthesize(texts, basenames, synth_dir, None, mel_filenames)
            for elems in zip(wav_filenames, mel_filenames, mel_output_filenames, speaker_ids, texts):
                file.write('|'.join([str(x) for x in elems]) + '\n')
    _merge_maps(synth_dir, rank, world)
    log('Predicted mel spectrograms are saved in {}'.format(synth_dir))
    return os.path.join(synth_dir, 'map.txt')


def _merge_maps(out_dir, rank, world):
    if world == 1:
        return
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()
    if rank == 0:
        with open(os.path.join(out_dir, 'map.txt'), 'w', encoding='utf-8') as out:
            for r in range(world):
                p = os.path.join(out_dir, 'map.rank%d.txt' % r)
                if os.path.exists(p):
                    out.write(open(p, encoding='utf-8').read())


def hparams_debug():
    from ..Utils.Hyperparams import hparams_debug_string
    return hparams_debug_string()


def tacotron_synthesize(args, hparams, checkpoint):
    return run_synthesis(args, _find_checkpoint(checkpoint), args.output_dir, hparams)


def tacotron_inference(args, hparams, checkpoint, sentences):
    if sentences is None:
        raise RuntimeError('Inference mode requires input sentence(s), make sure you put sentences in sentences.txt!')
    return run_inference(_find_checkpoint(checkpoint), args.output_dir, hparams, sentences)


def get_arguments(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument('--checkpoint', default='Tacotron_trained_logs/taco_pretrained/', help='Path to model checkpoint')
    parser.add_argument('--hparams', default='', help='Hyperparameter overrides as a comma-separated list of name=value pairs')
    parser.add_argument('--name', help='Name of logging directory if the two modules were trained together.')
    parser.add_argument('--tacotron_name', help='Name of logging directory of Tacotron. If trained separately')
    parser.add_argument('--model', default='Tacotron-2')
    parser.add_argument('--mode', default='synthesize', help='runing mode, could be synthesize or inference')
    parser.add_argument('--input_dir', default='Tacotron_input/', help='folder to contain inputs sentences/targets')
    parser.add_argument('--output_dir', default='tacotron_output/', help='folder to contain synthesized mel spectrograms')
    parser.add_argument('--GTA', default='True', help='Ground truth aligned synthesis, defaults to True, only considered in synthesis mode')
    parser.add_argument('--speaker_id', default=None, help='speaker ids list, comma separated')
    return parser.parse_args(argv)


def get_sentences(hparams, path='sentences.txt'):
    """sentences.txt of the working directory, else hparams.sentences (:129-141)."""
    sentences = []
    if os.path.exists(path):
        with open(path, 'rb') as f:
            sentences = [l.decode('utf-8').rstrip('\r\n') for l in f.readlines()]
        sentences = [s for s in sentences if s]
    return sentences if sentences else (hparams.sentences or None)


def main(argv=None):
    args = get_arguments(argv)
    modified_hparams = hparams.parse(args.hparams)
    if args.mode == 'synthesize':
        return tacotron_synthesize(args, modified_hparams, args.checkpoint)
    if args.mode == 'inference':
        return tacotron_inference(args, modified_hparams, args.checkpoint, get_sentences(modified_hparams))
    raise ValueError('unknown --mode %r (synthesize | inference)' % args.mode)


if __name__ == '__main__':
    main()
