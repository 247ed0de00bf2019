import glob

from distributed_resnet_tensorflow_b200.utils.events import EventFileWriter, read_events


def test_scalar_events_roundtrip(tmp_path):
    w = EventFileWriter(str(tmp_path))
    w.add_scalars({"cross_entropy": 2.5, "cost": 3.25, "learning_rate": 0.1, "Precision": 0.125}, step=100)
    w.add_scalars({"Best Precision": 0.5}, step=200)
    w.close()
    files = glob.glob(str(tmp_path / "events.out.tfevents.*"))
    assert len(files) == 1
    ev = list(read_events(files[0]))
    assert ev[0][1] == 0 and ev[0][2] == {}                      # file_version record
    assert ev[1][1] == 100 and abs(ev[1][2]["cost"] - 3.25) < 1e-6 and abs(ev[1][2]["Precision"] - 0.125) < 1e-6
    assert ev[2][1] == 200 and ev[2][2] == {"Best Precision": 0.5}


def test_events_readable_by_tensorboard_if_installed(tmp_path):
    import pytest
    ea = pytest.importorskip("tensorboard.backend.event_processing.event_accumulator")
    w = EventFileWriter(str(tmp_path))
    for s in range(5):
        w.add_scalars({"cost": float(s)}, step=s)
    w.close()
    acc = ea.EventAccumulator(str(tmp_path))
    acc.Reload()
    vals = [e.value for e in acc.Scalars("cost")]
    assert vals == [0.0, 1.0, 2.0, 3.0, 4.0]
