"""TEST INFRASTRUCTURE: an oracle-backed stand-in for `ep.CudaEPOps`, so the expert-parallel choreography (buffer layout,
fixed-capacity splits, slot maps, the three all-to-alls) can run on CPU tensors over `gloo`.  Each method restates the
contract of the C-ABI entry point of the same role (include/nf4moe.h) in numpy; the expert arithmetic is the oracle's."""
import numpy as np
import torch

from oracle import nf4_oracle as orc


class OracleEPOps:
    def __init__(self, dense13: np.ndarray, dense2: np.ndarray):
        self.w13, self.w2 = dense13, dense2    # this rank's experts, dequantised by the oracle: [E_local, 2I, H], [E_local, H, I]

    def route(self, logits, k):
        ids, w = orc.route_topk(logits.numpy(), k)
        return torch.from_numpy(ids), torch.from_numpy(w)

    def bucket(self, ids, E):                                       # nf4moe_bucket: stable counting sort by expert
        flat = ids.numpy().reshape(-1)
        keep = np.nonzero((flat >= 0) & (flat < E))[0]
        order = keep[np.argsort(flat[keep], kind="stable")]
        counts = np.bincount(flat[keep], minlength=E)
        offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
        return {"expert_offsets": torch.from_numpy(offs), "sorted_pairs": torch.from_numpy(order.astype(np.int32))}

    def pack(self, x, ids, b, E, world, cap, send_rows, send_ids, pair_slot, status):    # nf4moe_ep_pack
        k = ids.shape[1]
        el = E // world
        offs, sp = b["expert_offsets"].numpy(), b["sorted_pairs"].numpy()
        flat = ids.numpy().reshape(-1)
        pair_slot[: flat.size] = -1
        send_ids.fill_(-1)
        for p in range(world):
            start, cnt = offs[p * el], offs[(p + 1) * el] - offs[p * el]
            if cnt > cap:
                status[0] |= 1
            for j in range(min(cnt, cap)):
                pair = int(sp[start + j])
                send_rows[p * cap + j] = x[pair // k]
                send_ids[p * cap + j] = int(flat[pair]) - p * el
                pair_slot[pair] = p * cap + j

    def expert_rows(self, rows, row_expert, w13, w2, y_rows, regime):                     # nf4moe_expert_rows
        re = row_expert.numpy()
        r16 = rows.numpy()
        for e in range(self.w13.shape[0]):
            idx = np.nonzero(re == e)[0]
            if idx.size == 0:
                continue
            h = orc._mm_f16_f32acc(r16[idx], self.w13[e]).astype(np.float16)
            a = orc.swiglu_oai(h.astype(np.float32)).astype(np.float16)
            y = orc._mm_f16_f32acc(a, self.w2[e]).astype(np.float16)
            y_rows[torch.from_numpy(idx)] = torch.from_numpy(y)

    def combine(self, back_rows, weights, pair_slot, out):                                 # nf4moe_combine_indexed
        T, k = weights.shape
        y = back_rows.numpy().astype(np.float32)
        ps = pair_slot.numpy().reshape(T, k)
        w = weights.numpy()
        acc = np.zeros((T, y.shape[1]), dtype=np.float32)
        for s in range(k):
            ok = ps[:, s] >= 0
            acc[ok] += (w[ok, s:s + 1] * y[ps[ok, s]]).astype(np.float32)
        res = torch.from_numpy(acc.astype(np.float16))
        if out is not None:
            out.copy_(res)
            return out
        return res

    def local_forward(self, x, local_ids, weights, out):              # nf4moe_moe_forward with masked (-1) non-local pairs
        T, k = local_ids.shape
        ids, w, x16 = local_ids.numpy(), weights.numpy(), x.numpy()
        acc = np.zeros((T, self.w2.shape[1]), dtype=np.float32)
        for s in range(k):
            for e in range(self.w13.shape[0]):
                idx = np.nonzero(ids[:, s] == e)[0]
                if idx.size == 0:
                    continue
                h = orc._mm_f16_f32acc(x16[idx], self.w13[e]).astype(np.float16)
                a = orc.swiglu_oai(h.astype(np.float32)).astype(np.float16)
                y = orc._mm_f16_f32acc(a, self.w2[e]).astype(np.float16)
                acc[idx] += (w[idx, s:s + 1] * y.astype(np.float32)).astype(np.float32)
        res = torch.from_numpy(acc.astype(np.float16))
        if out is not None:
            out.copy_(res)
            return out
        return res
