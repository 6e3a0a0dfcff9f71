"""The drop-in boundary: ChatOpenAI-compatible constructor, Runnable
protocol, batching across concurrent ainvoke calls, cancellation and the error
contract of app.py:177-197 -- exercised with a stub engine on the CPU, and inside
the reference's own unmodified app.py when it is present."""
import asyncio
import threading
import time

import pytest

import kubectl_b200  # noqa: F401
from kubectl_b200 import synth
from kubectl_b200.chat_model import B200ChatModel, CallableEngine, Engine, prompt_to_text
from oracle import ref_loader


def make(fn=lambda p: "kubectl get pods", **kw):
    eng = CallableEngine(fn)
    return B200ChatModel(temperature=0, api_key="x", model="gpt-3.5-turbo", request_timeout=60, engine=eng, **kw), eng


def test_constructor_takes_the_reference_kwargs():
    m, _ = make(base_url="http://ignored")
    assert m.arch == "qwen2.5-0.5b"          # unknown OpenAI name -> default local architecture
    m.close()
    m2 = B200ChatModel(temperature=0, api_key=None, model="Llama-3.2-1B", request_timeout=5, engine=CallableEngine(str))
    assert m2.arch == "llama-3.2-1b"
    m2.close()
    with pytest.raises(ValueError):
        B200ChatModel(temperature=0.7, model="x", engine=CallableEngine(str))


def test_invoke_batch_and_pipe():
    m, eng = make(lambda p: "echo:" + p)
    assert m.invoke("hi").content == "echo:hi"
    assert [x.content for x in m.batch(["a", "b", "c"])] == ["echo:a", "echo:b", "echo:c"]

    class Tmpl:
        def invoke(self, v, config=None):
            return "Q=" + v["query"]

    class Parser:
        def invoke(self, v, config=None):
            return prompt_to_text(v).upper()

    chain = Tmpl() | m | Parser()
    assert chain.invoke({"query": "x"}) == "ECHO:Q=X"
    assert asyncio.run(chain.ainvoke({"query": "y"})) == "ECHO:Q=Y"
    m.close()


def test_concurrent_ainvoke_calls_are_batched():
    m, eng = make(lambda p: p[::-1], max_wait_ms=50)

    async def go():
        return await asyncio.gather(*(m.ainvoke(f"q{i}") for i in range(40)))
    outs = asyncio.run(go())
    assert [o.content for o in outs] == [f"q{i}"[::-1] for i in range(40)]
    assert max(eng.batches) > 1 and sum(eng.batches) == 40
    m.close()


def test_wait_for_timeout_cancels_and_request_is_dropped():
    gate = threading.Event()

    class Slow(Engine):
        max_batch = 1
        seen = []

        def generate(self, prompts, cancelled=None):
            self.seen.extend(prompts)
            gate.wait(2)
            return ["kubectl get pods"] * len(prompts)

    eng = Slow()
    m = B200ChatModel(temperature=0, model="x", request_timeout=60, engine=eng, max_wait_ms=1)

    async def go():
        t1 = asyncio.ensure_future(m.ainvoke("first"))       # occupies the engine
        await asyncio.sleep(0.05)
        with pytest.raises(asyncio.TimeoutError):
            await asyncio.wait_for(m.ainvoke("second"), timeout=0.05)   # app.py:183-186 -> 504
        gate.set()
        return (await t1).content
    assert asyncio.run(go()) == "kubectl get pods"
    time.sleep(0.1)
    assert "second" not in eng.seen and m.frontend.dropped == 1
    m.close()


def test_engine_failure_surfaces_as_runtime_error_not_value_error():
    def boom(p):
        raise KeyError("kv pages exhausted")
    m, _ = make(boom)
    with pytest.raises(RuntimeError, match="kv pages exhausted"):
        m.invoke("x")

    async def go():
        await m.ainvoke("y")
    with pytest.raises(RuntimeError):
        asyncio.run(go())
    m.close()


def test_one_failed_request_does_not_fail_its_batch():
    """An engine may return an exception instance for one prompt (e.g. a prompt that does not fit its context): that
    request gets a RuntimeError (-> 500, app.py:195-197), the others of the same batch their completions."""
    m, eng = make(lambda p: OverflowError("prompt too long") if "LONG" in p else "kubectl get pods", max_wait_ms=50)
    outs = [None, None, None]

    def call(i, p):
        try:
            outs[i] = m.invoke(p).content
        except BaseException as e:  # noqa: BLE001
            outs[i] = e
    ts = [threading.Thread(target=call, args=(i, p)) for i, p in enumerate(["a", "LONG b", "c"])]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert outs[0] == outs[2] == "kubectl get pods"
    assert isinstance(outs[1], RuntimeError) and not isinstance(outs[1], ValueError) and "prompt too long" in str(outs[1])
    assert max(eng.batches) >= 2                                # they really shared a batch
    m.close()


@pytest.mark.skipif(not ref_loader.reference_available(), reason="/root/reference only exists in the build container")
def test_zero_diff_install_into_unmodified_app_py():
    """app.py is executed verbatim; only `langchain_openai.ChatOpenAI` is ours."""
    from starlette.testclient import TestClient
    made = []

    class Local(B200ChatModel):
        def __init__(self, **kw):
            super().__init__(engine=CallableEngine(lambda p: synth.stub_completion(p, unsafe_rate=0.0)), **kw)
            made.append(self)

    app = ref_loader.load_reference_app(chat_openai=Local, module_name="kx_ref_app_boundary")
    assert app.chain is not None and len(made) == 1
    assert made[0].request_timeout == app.LLM_TIMEOUT and made[0].model_name == app.OPENAI_MODEL
    with TestClient(app.app) as c:
        r = c.post("/kubectl-command", json={"query": "list all pods in namespace ns7"})
        assert r.status_code == 200 and r.json()["from_cache"] is False
        cmd = r.json()["kubectl_command"]
        assert cmd.startswith("kubectl get ")
        r = c.post("/kubectl-command", json={"query": "  list all\tpods in namespace   ns7\n"})
        assert r.status_code == 200 and r.json()["from_cache"] is True and r.json()["kubectl_command"] == cmd
        # unsafe completion -> parser raises ValueError -> 422 (app.py:192-194)
        made[0].engine.fn = lambda p: "kubectl get pods; rm -rf /"
        r = c.post("/kubectl-command", json={"query": "something new"})
        assert r.status_code == 422 and "LLM generated unsafe command" in r.json()["detail"]
        # engine failure -> 500 with the message in the detail (app.py:195-197)
        def boom(p):
            raise MemoryError("out of KV pages")
        made[0].engine.fn = boom
        r = c.post("/kubectl-command", json={"query": "another new one"})
        assert r.status_code == 500 and "out of KV pages" in r.json()["detail"]
        assert c.post("/kubectl-command", json={"query": "ab"}).status_code == 422    # pydantic min_length
    made[0].close()


def test_identical_prompts_in_flight_are_generated_once():
    m, eng = make(lambda p: p.upper(), max_wait_ms=50)

    async def go():
        return await asyncio.gather(*(m.ainvoke(f"q{i % 3}") for i in range(30)))
    outs = asyncio.run(go())
    assert [o.content for o in outs] == [f"Q{i % 3}" for i in range(30)]
    assert sum(eng.batches) < 30 and m.frontend.coalesced == 30 - sum(eng.batches)
    m.close()


def test_chat_model_hooks_generate_and_agenerate():
    """The BaseChatModel hooks ChatOpenAI implements (_generate / _agenerate / _llm_type / _identifying_params); LangChain
    itself is not in this image, so the result objects are the plain stand-ins with the same attributes."""
    m, _ = make(lambda p: "kubectl get pods\nExplanation: lists pods")

    class Msg:
        def __init__(self, content):
            self.content = content

    res = m._generate([Msg("system text"), Msg("list pods")], stop=["\nExplanation"])
    assert res.generations[0].message.content == "kubectl get pods" and res.llm_output["model_name"] == m.arch
    res2 = asyncio.run(m._agenerate([Msg("list pods")]))
    assert res2.generations[0].message.content.startswith("kubectl get pods\n")
    assert m._llm_type == "b200-local" and m._identifying_params == {"model_name": m.arch, "temperature": 0}
    try:
        import langchain_core  # noqa: F401
    except ImportError:
        with pytest.raises(ImportError):
            m.langchain_chat_model()
    m.close()
