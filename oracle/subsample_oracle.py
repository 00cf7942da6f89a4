"""TEST INFRASTRUCTURE ONLY: CPU restatement of masked_FiT's ImageNetLatentIterator.subsample + collate stacking
(masked_FiT/preprocess/iterators.py:142-169) with torch on CPU.  The reference class needs its dataset files to be constructed, so the
two methods are restated (12 lines); the RNG calls (torch.randperm per over-long sample, in batch order) are the reference's.
PINNED: oracle/make_goldens.py::gen_masked calls the reference's own _patchify / subsample / collate unbound (object made with __new__)
and tests/test_oracle_golden.py::test_masked_fit_collate_and_training_step_match_reference checks this restatement against that output
bit for bit (tests/golden/masked_fit_tiny.npz)."""
import torch


def subsample(latent, pos, mask, number_of_tokens, P, embed_dim, generator=None):
    if latent.shape[0] > number_of_tokens:                                      # iterators.py:143
        permutation = torch.randperm(latent.shape[0], generator=generator)       # :144
        latent = latent[permutation, :][:number_of_tokens, :]                    # :146
        pos = pos[permutation, :][:number_of_tokens, :]                          # :147
        mask = mask[permutation][:number_of_tokens]                              # :148
    else:
        latent = torch.nn.functional.pad(latent, (0, P - latent.shape[1], 0, number_of_tokens - latent.shape[0]))    # :150-151
        pos = torch.nn.functional.pad(pos, (0, embed_dim - pos.shape[1], 0, number_of_tokens - pos.shape[0]))        # :152
        mask = torch.nn.functional.pad(mask, (0, number_of_tokens - mask.shape[0]))                                  # :153
    return latent, pos, mask


def collate(latents, poss, masks, number_of_tokens, generator=None):
    P, d = latents[0].shape[1], poss[0].shape[1]
    out = [subsample(l, p, m, number_of_tokens, P, d, generator) for l, p, m in zip(latents, poss, masks)]     # :160
    lat, pos, msk = zip(*out)
    return torch.stack(lat), torch.stack(pos), torch.stack(msk)                                                # :164-167
