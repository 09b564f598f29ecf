"""Filter URL lists: drop blacklisted domains, non-text extensions, malformed and duplicate URLs.
usage: python blacklist_urls.py <directory of url files> <output file>"""
import glob
import re
import sys
import time
from urllib.parse import urlparse

DOMAIN_BLACKLIST = {
    "500px", "aapks", "akamaihd", "amazon", "apple", "artifactfire", "artstation", "awwni", "bandcamp", "battleforthenet",
    "coinscalendar", "dailymotion", "deviantart", "discord", "discordapp", "dlapkandroid", "dropbox", "e621", "ebay",
    "edealinfo", "erome", "eroshare", "explosm", "facebook", "fbcdn", "flickr", "furaffinity", "futhead", "gatopardo",
    "gfycat", "gifsound", "gifsoup", "giphy", "github", "google", "gunprime", "gyazo", "hotdealstar", "imagefap",
    "imageshack", "imgflip", "imgur", "instagram", "karmadecay", "kryptocal", "kym-cdn", "liveleak", "livememe",
    "lmgtfy", "magaimg", "memegenerator", "minorplanetcenter", "minus", "mobafire", "morejpeg", "nocookie",
    "pcpartpicker", "photobucket", "pinimg", "pinterest", "pixiv", "pornhub", "prntscr", "puu", "qkme", "quickmeme",
    "radd", "redd", "reddit", "reddit-stream", "redditlog", "redditmedia", "reddituploads", "redtube", "reupp",
    "reverb", "roanoke", "rollingstone", "sli", "soundcloud", "soundgasm", "spankbang", "spotify", "strawpoll",
    "streamable", "timeanddate", "tinypic", "touhouradio", "tumblr", "twimg", "twitch", "twitter", "vid", "vimeo",
    "vine", "vkaao", "vocaroo", "voyagefusion", "walmart", "wciu", "wikimedia", "wikipedia", "xhamster", "xkcd",
    "xvideos", "youtu", "youtube", "youtubedoubler", "ytimg", "zillexplorer",
}
EXTENSION_BLACKLIST = (".3gp", ".7z", ".ai", ".aif", ".apk", ".app", ".avi", ".bin", ".bmp", ".bz2", ".css", ".csv",
                       ".dat", ".deb", ".dmg", ".doc", ".docx", ".exe", ".gif", ".gifv", ".gz", ".iso", ".jar", ".jpeg",
                       ".jpg", ".js", ".log", ".mid", ".midi", ".mkv", ".mov", ".mp3", ".mp4", ".mpeg", ".mpg", ".ogg",
                       ".ogv", ".otf", ".pdf", ".pkg", ".png", ".pps", ".ppt", ".pptx", ".psd", ".py", ".qt", ".ram",
                       ".rar", ".sql", ".svg", ".swf", ".tar.gz", ".tar", ".tgz", ".tiff", ".ttf", ".txt", ".wav",
                       ".webm", ".wma", ".wmv", ".xls", ".xlsx", ".xml", ".xz", ".zip")
URL_RE = re.compile(r"^(?:http)s?://(?:(?:[A-Z0-9](?:[A-Z0-9-]{0,61}[A-Z0-9])?\.)+(?:[A-Z]{2,6}\.?|[A-Z0-9-]{2,}\.?)|"
                    r"localhost|\d{1,3}\.\d{1,3}\.\d{1,3}\.\d{1,3})(?::\d+)?(?:/?|[/?]\S+)$", re.IGNORECASE)


def domain_of(url):
    try:
        import tldextract
        return tldextract.extract(url).domain
    except ImportError:
        host = urlparse(url).netloc.split(":")[0].lower().split(".")
        return host[-2] if len(host) >= 2 else (host[0] if host else "")


def is_bad(url):
    if domain_of(url) in DOMAIN_BLACKLIST:
        return "domain"
    path = url.split("?")[0].lower()
    if path.endswith(EXTENSION_BLACKLIST):
        return "extension"
    if len(url) <= 8:
        return "short"
    if URL_RE.match(url) is None:
        return "malformed"
    return None


if __name__ == "__main__":
    t0 = time.time()
    if len(sys.argv) < 3 or sys.argv[1] in ("-h", "--help"):
        sys.exit("usage: python tools/openwebtext/blacklist_urls.py <directory with url files> <output file of clean urls>")
    path, out = sys.argv[1], sys.argv[2]
    seen, counts = set(), {}
    for fname in sorted(glob.glob(path + "/*.txt")):
        with open(fname, encoding="utf-8", errors="ignore") as f:
            for line in f:
                url = line.strip()
                why = is_bad(url) or ("duplicate" if url in seen else None)
                if why:
                    counts[why] = counts.get(why, 0) + 1
                    continue
                seen.add(url)
    with open(out, "w") as f:
        for url in seen:
            f.write(url + "\n")
    print(f"kept {len(seen)} urls, dropped {counts} in {time.time() - t0:.1f} s")
