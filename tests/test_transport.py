import threading

import torch

from fedrec_b200.utils.transport import StateDictServer, average_state_dicts, send_state_dict


def test_tcp_upload_and_average(tmp_path):
    srv = StateDictServer("127.0.0.1", 0, out_dir=str(tmp_path))
    sds = [{"w": torch.full((3, 4), float(i + 1)), "n": torch.tensor(7)} for i in range(3)]
    out = {}
    t = threading.Thread(target=lambda: out.setdefault("paths", srv.receive_round(3, timeout_s=30)))
    t.start()
    for i, sd in enumerate(sds):
        assert send_state_dict(sd, "127.0.0.1", srv.port, client_id=i) > 0
    t.join(30)
    srv.close()
    assert len(out["paths"]) == 3
    avg = average_state_dicts(out["paths"])
    torch.testing.assert_close(avg["w"], torch.full((3, 4), 2.0))
    assert int(avg["n"]) == 7
    wavg = average_state_dicts(sorted(out["paths"]), weights=[1, 0, 0])
    assert float(wavg["w"][0, 0]) in (1.0, 2.0, 3.0)
