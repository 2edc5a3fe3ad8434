"""Writes tests/golden/reference_doc_samples.json: the TaskModel JSON bodies the reference's own tutorial
text shows (run in the build container, where /root/reference is mounted; the GPU box only reads the
committed JSON).  Each entry keeps the text exactly as the tutorial prints it (indentation and member
order included) and the member values a reader of System.Text.Json's documentation expects from it."""
import json
import os
import re

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

# (file, first line, last line) of each fenced JSON body that is a TaskModel (or its "data" member)
SOURCES = [
    ("docs/aca/05-aca-dapr-pubsubapi/index.md", 61, 68),      # body published to tasksavedtopic
    ("docs/aca/09-aca-autoscale-keda/index.md", 196, 205),    # the "data" object of the Service Bus message
    ("docs/aca/03-aca-dapr-integration/index.md", 120, 125),  # POST api/tasks body (TaskAddModel members)
    ("docs/aca/06-aca-dapr-bindingsapi/index.md", 270, 275),  # queue message handled by ExternalTasksProcessor
]


def main():
    out = []
    for path, a, b in SOURCES:
        lines = open(os.path.join(REF, path), encoding="utf-8").read().split("\n")[a - 1:b]
        text = "\n".join(lines)
        if '"data"' in text.split("\n")[0] or text.lstrip().startswith('"data"'):
            text = text[text.index("{"):]
        text = text.strip()
        if not text.startswith("{"):
            text = "{" + text
        if text.count("{") > text.count("}"):
            text += "}"
        doc = json.loads(text)  # what any JSON reader sees; TaskModel members only
        out.append({"source": f"{path}:{a}-{b}", "text": text, "members": doc})
    with open(os.path.join(HERE, "reference_doc_samples.json"), "w", encoding="utf-8") as f:
        json.dump(out, f, indent=1, ensure_ascii=True)
    for e in out:
        print(e["source"], sorted(e["members"]))


if __name__ == "__main__":
    main()
