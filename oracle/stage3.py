"""CPU restatement of the reference's stage 3 shard writer (TEST INFRASTRUCTURE ONLY; row N1 of SURVEY.md 8(f)).

Only tests/ and bench.py's CPU legs import this.  Follows tools/make_parquet_list.py:26-62 (`job`) and :100-108 (the
three torch.load calls): Python-list dicts -> a pandas frame of Python lists -> DataFrame.to_parquet.
"""
import pandas as pd
import torch


def load_pt(src_dir):
    """make_parquet_list.py:100-102."""
    return (torch.load(f"{src_dir}/utt2embedding.pt"), torch.load(f"{src_dir}/spk2embedding.pt"),
            torch.load(f"{src_dir}/utt2speech_token.pt"))


def job(utt_list, parquet_file, utt2wav, utt2text, utt2spk, utt2embedding, spk2embedding, utt2speech_token,
        utt2instruct=None):
    """make_parquet_list.py:26-57 (without the tqdm bar and the two json side files)."""
    df = pd.DataFrame()
    df['utt'] = utt_list
    df['wav'] = [utt2wav[u] for u in utt_list]
    df['audio_data'] = [open(utt2wav[u], 'rb').read() for u in utt_list]
    df['text'] = [utt2text[u] for u in utt_list]
    df['spk'] = [utt2spk[u] for u in utt_list]
    df['utt_embedding'] = [utt2embedding[u] for u in utt_list]
    df['spk_embedding'] = [spk2embedding[utt2spk[u]] for u in utt_list]
    df['speech_token'] = [utt2speech_token.get(u, []) for u in utt_list]
    if utt2instruct is not None:
        df['instruct'] = [utt2instruct[u] for u in utt_list]
    df.to_parquet(parquet_file)
