"""Recorded-response stand-in for reference common/openai_generic_assistant.py:10-135.

There is no network here, so the assistant answers from a table of recorded responses: the first
entry whose key occurs in the last message wins, else the default.  A response may be a LIST: the
answers to successive prompts that match the key, the last one repeating - a recorded first answer
that is malformed followed by the corrected one is how the retry loops of the reference's drivers
(test_all.py:63-83, :99-122) are replayed.  The method names and the shape of what
wait_get_last_k_message returns (`.data[0].content[0].text.value`) are the reference's; polling
never sleeps.

The reference's set-up functions construct the assistant without arguments and name it in
create_assistant (find_srckind_metapath_neo4j.py:46-50, generate_query.py:18-23,
analyze_root_cause.py:6-12); `install_recorded(name, table)` registers the table such an assistant
picks up under that name."""
import time
from types import SimpleNamespace

_BY_NAME: dict = {}


def install_recorded(name, table, default=None):
    """Recorded answers for assistants created under `name` without a table of their own."""
    _BY_NAME[name] = (dict(table), default)


def _message(text):
    return SimpleNamespace(content=[SimpleNamespace(text=SimpleNamespace(value=text))], role="assistant")


class OpenAIGenericAssistant:
    def __init__(self, recorded=None, default="(recorded assistant: no response recorded for this prompt)"):
        self._own_table = recorded is not None
        self.recorded = list((recorded or {}).items())
        self.default = default
        self.assistant = self.thread = self.run = None
        self.messages = []          # newest first, like the API's list order
        self.runs = []
        self._served = {}           # key -> how many prompts it has answered

    def create_assistant(self, instructions, name, model="gpt-4"):
        self.assistant = SimpleNamespace(id="asst_recorded_" + str(name), instructions=instructions, name=name, model=model)
        if not self._own_table and name in _BY_NAME:
            table, default = _BY_NAME[name]
            self.recorded = list(table.items())
            if default is not None:
                self.default = default

    def retrieve_assistant(self, assistant_id):
        self.assistant = SimpleNamespace(id=assistant_id)

    def create_thread(self):
        self.thread = SimpleNamespace(id="thread_recorded")

    def retrieve_thread(self, thread_id):
        self.thread = SimpleNamespace(id=thread_id)

    def add_message(self, content):
        self.messages.insert(0, SimpleNamespace(content=[SimpleNamespace(text=SimpleNamespace(value=content))], role="user"))

    def run_assistant(self, instructions=None):
        prompt = next((m.content[0].text.value for m in self.messages if m.role == "user"), "")
        answer = self.default
        for key, resp in self.recorded:
            if key in prompt:
                if isinstance(resp, (list, tuple)):
                    k = self._served.get(key, 0)
                    self._served[key] = k + 1
                    resp = resp[min(k, len(resp) - 1)]
                answer = resp
                break
        now = int(time.time())
        self.run = SimpleNamespace(id="run_%d" % len(self.runs), status="completed", created_at=now, completed_at=now,
                                   usage=SimpleNamespace(prompt_tokens=len(prompt.split()),
                                                         completion_tokens=len(answer.split()),
                                                         total_tokens=len(prompt.split()) + len(answer.split())))
        self.runs.insert(0, self.run)
        self.messages.insert(0, _message(answer))

    def get_run_status(self):
        return self.run

    def display_response(self):
        for m in reversed(self.messages):
            print(m.role + ": " + m.content[0].text.value)

    def get_last_message(self):
        return SimpleNamespace(data=self.messages[:1])

    def get_all_message(self):
        return SimpleNamespace(data=list(self.messages))

    def get_last_k_message(self, num):
        return SimpleNamespace(data=self.messages[:num])

    def wait_get_last_k_message(self, num=1):
        return self.get_last_k_message(num)

    def get_token_usage(self, tmin, tmax, limit=20):
        used = [r.usage for r in self.runs[:limit]
                if tmin <= r.created_at < tmax and tmin <= r.completed_at < tmax]
        return {"prompt_tokens": sum(u.prompt_tokens for u in used),
                "completion_tokens": sum(u.completion_tokens for u in used),
                "total_tokens": sum(u.total_tokens for u in used)}
