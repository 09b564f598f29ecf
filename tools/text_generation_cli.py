"""Interactive client for the generation server (tools/text_generation_cli.py:1-23; python 3).

    python tools/text_generation_cli.py localhost:5000
"""
import json
import sys
import urllib.request


def put(url, payload):
    req = urllib.request.Request(url, data=json.dumps(payload).encode(), method="PUT",
                                 headers={"Content-Type": "application/json; charset=UTF-8"})
    with urllib.request.urlopen(req) as r:
        return json.loads(r.read().decode())


def main():
    if len(sys.argv) < 2:
        raise SystemExit("usage: text_generation_cli.py host:port")
    url = "http://" + sys.argv[1] + "/api"
    while True:
        try:
            sentence = input("Enter prompt: ")
            n = int(input("Enter number of tokens to generate: "))
        except (EOFError, KeyboardInterrupt):
            print()
            return
        out = put(url, {"prompts": [sentence], "tokens_to_generate": n})
        print("Megatron Response: ")
        print(out["text"][0] if isinstance(out, dict) else out)


if __name__ == "__main__":
    main()
