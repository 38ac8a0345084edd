"""Synthetic vocab and texts shared by the tokenizer tests (no published vocab.txt in this image)."""
VOCAB = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]", "hello", "world", "##s", "what", "is", "the", "un",
         "##believ", "##able", "!", "?", ",", ".", "a", "b", "c", "##a", "##b", "##c", "cafe", "e", "##e",
         "explain", "how", "to", "write", "code", "in", "python", "##ic", "-", "'", "don", "t", "1", "##2", "##3",
         "(", ")", "new", "york", "##er", "x", "##x", "中", "文", "$", "def", "_", "foo", "##_", "##foo"]

TEXTS = [
    "Hello worlds! What is the unbelievable cafe abc?",
    "  explain   how to write\tpythonic code\n in python... ",
    "don't (new-yorker) 123 xxx $x",
    "Café ÉE 中文abc",
    "",
    "zzzz unknownword hello",
    "def foo_foo(a, b): x_x",
    "a" * 120 + " hello",
    "hello " * 40,
]
