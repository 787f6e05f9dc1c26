# hparams for FastSpeech2 (plain Python assignments, loaded by utils.hparams.configure of the
# reference, utils/__init__.py:24-46).  The reference README names config/hparams_template.py but
# does not ship it; values follow utils/default.yaml with two documented departures:
#   reduction_rate = 1          (the yaml's 2 is for Transformer-TTS; datasets force 1 for FastSpeech2)
#   energy_min/max = 0.0/403.8  (the yaml has them swapped, which would give descending bins)
architecture = 'text-mel'
model = 'Fastspeech2'
vocab_size = 152
mel_dim = 80
amp = True

# data (placeholders: the benchmarks and tests use synthetic batches)
train_script = 'data/train.csv'
test_script = 'data/test.csv'
lengths_file = 'data/lengths.npy'
save_dir = 'checkpoints/fastspeech2'
mean_file = None
var_file = None
spm_model = None
different_spk_emb_samespeaker = False
loaded_epoch = None
loaded_dir = None

# optimisation
optimizer = 'Noam'
warmup_step = 4000
warmup_factor = 1.0
max_seqlen = 10000
batch_size = None
max_epoch = 200
save_per_epoch = 50
clip = 1.0

# encoder
d_model_encoder = 384
n_layer_encoder = 6
n_head_encoder = 4
ff_conv_kernel_size_encoder = 5
concat_after_encoder = False

# decoder
d_model_decoder = 384
n_layer_decoder = 6
n_head_decoder = 4
ff_conv_kernel_size_decoder = 1
concat_after_decoder = False

postnet_pred = True
reduction_rate = 1
dropout = 0.1

# variance adaptor
nbins = 256
f0_min = 71.0
f0_max = 799.8
energy_min = 0.0
energy_max = 403.8
