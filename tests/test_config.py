"""configs/*.yaml schema (alegpu_config_*; mirror of load_config, src/bin/train.cc:108-136).

tests/golden/configs/{debug,v0,v1}.yaml are byte-for-byte copies of the reference's shipped configs.
No GPU needed: the loader is host code inside libalegpu.so."""
import os
import shutil

import pytest

import alegpu

HERE = os.path.dirname(os.path.abspath(__file__))
CFG = os.path.join(HERE, "golden", "configs")

# (file) -> the values read off /root/reference/configs/<file>
EXPECT = {
    "debug": dict(total_environments=8, hidden_size=32, horizon=32, max_steps=64800, frame_stack=4,
                  learning_rate=2.5e-4, clip_param=0.2, value_loss_coef=0.5, entropy_coef=0.01,
                  num_epochs=2, mini_batch_size=64, num_mini_batches=4, gae_discount=0.99,
                  gae_lambda=0.95, max_gradient_norm=0.5, num_rollouts=10, num_workers=1,
                  worker_batch_size=1, frame_skip=4, max_return=864, record_observation=0,
                  record_video=1, cuda_graph=0, deterministic=1),
    "v0": dict(total_environments=8, hidden_size=512, horizon=128, learning_rate=0.00025,
               clip_param=0.1, value_loss_coef=0.5, num_epochs=4, mini_batch_size=256,
               num_mini_batches=4, num_rollouts=9760, num_workers=8, worker_batch_size=1),
    "v1": dict(total_environments=4096, hidden_size=512, horizon=5, clip_param=0.2,
               value_loss_coef=0.4, num_epochs=1, mini_batch_size=1280, num_mini_batches=16,
               num_rollouts=1000000, num_workers=32, worker_batch_size=8),
}


@pytest.mark.parametrize("name", sorted(EXPECT))
def test_shipped_configs_parse_like_the_reference(name):
    c = alegpu.Config.load(os.path.join(CFG, name + ".yaml"))
    for k, v in EXPECT[name].items():
        assert getattr(c, k) == pytest.approx(v, rel=1e-6), (k, getattr(c, k), v)
    assert c.action_size == 4  # the YAML key is never read (src/bin/train.cc:36)
    assert c.mini_batch_size * c.num_mini_batches == c.total_environments * c.horizon  # D6
    hp = c.hyperparameters()  # prepare_hyperparameters, src/bin/train.cc:285-291
    assert hp.clip_param == pytest.approx(EXPECT[name]["clip_param"])


def test_copies_equal_the_reference_files():
    ref = "/root/reference/configs"
    if not os.path.isdir(ref):
        pytest.skip("reference tree not present (GPU box)")
    for n in EXPECT:
        with open(os.path.join(ref, n + ".yaml"), "rb") as a, open(os.path.join(CFG, n + ".yaml"), "rb") as b:
            assert a.read() == b.read()


def test_defaults_are_load_configs_inline_defaults():
    c = alegpu.Config.parse("", validate=False)  # every key absent: src/bin/train.cc:111-134
    want = dict(total_environments=512, hidden_size=512, horizon=128, max_steps=108000, frame_stack=4,
                learning_rate=2.5e-4, clip_param=0.1, value_loss_coef=0.5, entropy_coef=0.01,
                num_epochs=1, mini_batch_size=2048, num_mini_batches=32, gae_discount=0.99,
                gae_lambda=0.95, max_gradient_norm=0.5, num_rollouts=7000, num_workers=16,
                worker_batch_size=32, frame_skip=4, max_return=-1.0, record_observation=0,
                record_video=0, cuda_graph=0, deterministic=0, action_size=4,
                normalize_advantages=0, clip_value_loss=0, shuffle=0, num_gpus=1)
    for k, v in want.items():
        assert getattr(c, k) == pytest.approx(v), k
    assert set(want) == {k for k, _ in alegpu.Config._fields_}
    c.validate()  # 512 * 128 == 2048 * 32


def test_syntax_comments_quotes_unknown_keys_and_action_size():
    c = alegpu.Config.parse("""
# a comment
---
total_environments: 16   # trailing comment
horizon: "8"
mini_batch_size: 32
num_mini_batches: 4
learning_rate: 1e-3
deterministic: yes
record_video: Off
action_size: 18
some_future_key: [1, 2]
clip_value_loss: true
""")
    assert (c.total_environments, c.horizon, c.learning_rate) == (16, 8, 1e-3)
    assert c.deterministic == 1 and c.record_video == 0 and c.clip_value_loss == 1
    assert c.action_size == 4


def test_bad_file_and_bad_conversion_are_runtime_errors(tmp_path):
    with pytest.raises(RuntimeError):  # YAML::BadFile
        alegpu.Config.load(str(tmp_path / "missing.yaml"))
    for text in ("horizon: abc", "total_environments: -3", "deterministic: maybe", "clip_param: 0.1x",
                 "just a line"):
        with pytest.raises(RuntimeError):  # YAML::BadConversion / parser error
            alegpu.Config.parse(text, validate=False)


def test_d6_validation(tmp_path):
    shutil.copy(os.path.join(CFG, "v0.yaml"), tmp_path / "c.yaml")
    text = open(tmp_path / "c.yaml").read().replace("mini_batch_size: 256", "mini_batch_size: 128")
    with pytest.raises(ValueError):  # mini_batch_size * num_mini_batches != E * T
        alegpu.Config.parse(text)
    with pytest.raises(RuntimeError):  # train.h:140-143
        alegpu.Config.parse("total_environments: 3\nhorizon: 5\nnum_mini_batches: 2\nmini_batch_size: 7")
    with pytest.raises(ValueError):
        alegpu.Config.parse("frame_stack: 3")
    with pytest.raises(ValueError):  # 1024 / 4 = 256 rows per minibatch, not divisible by 3 ranks
        alegpu.Config.parse(text.replace("mini_batch_size: 128", "mini_batch_size: 256") + "\nnum_gpus: 3")
    assert alegpu.Config.parse(open(os.path.join(CFG, "v1.yaml")).read() + "\nnum_gpus: 8").num_gpus == 8
