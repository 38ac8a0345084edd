"""Batch entry at the orchestrator level: `route_queries(router, histories)`.

The reference's orchestrator takes ONE conversation per call (`Router.route_query(history)`,
src/router.py:199; `/chat` posts one message, src/app.py:27-28), so nothing above the routing engine
can feed the GPU a batch.  This module adds the missing sibling WITHOUT touching or copying
`router.py`: the routing decisions of all conversations are computed in one fused device pass
(`QueryRouter.route_queries`), and then the reference's own, unmodified `Router.route_query` is run
per conversation with that decision handed to it — response cache, device call, failover, perf
feedback and the result dict are all the reference's code.

    from router import Router                 # the reference's orchestrator (user's src/)
    from router_batch import route_queries    # this package
    results = route_queries(router, [history_a, history_b, ...])   # [(dict, tokens, device), ...]

Equals `[router.route_query(h) for h in histories]` (apart from `routing_overhead_ms`) whenever no
conversation of the batch depends on an earlier one of the same batch (same caveat as
`QueryRouter.route_queries`).  One batch at a time per Router object: the call swaps the router's
engine for a replaying proxy while it runs.
"""
from __future__ import annotations

import threading
from typing import Any, Dict, List, Sequence, Tuple

_LOCK = threading.Lock()


class _Replay:
    """Stands in for the QueryRouter during ONE `Router.route_query` call: hands out the decision the
    batch pass already made, forwards everything else (strategy name, update_perf, cache
    passthroughs ...) to the real engine."""

    def __init__(self, real, decision):
        self.__dict__["_real"], self.__dict__["_decision"] = real, decision

    def route_query(self, query=None, context=None, context_key=None):
        return self.__dict__["_decision"]

    def __getattr__(self, name):
        return getattr(self.__dict__["_real"], name)

    def __setattr__(self, name, value):
        setattr(self.__dict__["_real"], name, value)


def route_queries(router, histories: Sequence[List[Dict[str, Any]]]) -> List[Tuple[Dict[str, Any], int, str]]:
    """`router`: a reference `Router`; `histories`: one conversation history (list of
    {"role", "content"} dicts) per request.  Returns what `route_query` returns, per request."""
    with _LOCK:
        real = router.query_router
        prepared = [router._history_to_query_and_context(h) for h in histories]
        need = []
        for i, (query, _ctx, ctx_hash) in enumerate(prepared):
            if router.enable_response_cache and \
                    router._get_response_cache(router._response_cache_key(ctx_hash, query)) is not None:
                continue                      # answered from the response cache: no routing at all
            need.append(i)
        decisions: Dict[int, Any] = {}
        if need and hasattr(real, "route_queries"):
            try:
                ds = real.route_queries([prepared[i][0] for i in need], [prepared[i][1] for i in need],
                                        [prepared[i][2] for i in need])
                decisions = dict(zip(need, ds))
            except Exception:                 # noqa: BLE001 — the per-request path has its own fallback
                decisions = {}
        out = []
        try:
            for i, h in enumerate(histories):
                router.query_router = _Replay(real, decisions[i]) if i in decisions else real
                out.append(router.route_query(h))
        finally:
            router.query_router = real
        return out
